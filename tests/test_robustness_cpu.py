"""Malformed inputs end in an error message, not in a crash: structurally mutated network configs through the module / optimizer
factories, and mutated transforms.json files through the C++ loader's describer. (The byte-level decoders - JSON, msgpack, PNG,
JPEG, EXR, BMP, GIF, PSD, PIC, TGA - were run under ASan / UBSan on tens of thousands of mutated files while they were written;
tests/test_host_io_cpu.py keeps a few hundred of those cases.)"""
import copy
import ctypes
import json
import os
import random

import pytest

JUNK = [None, True, "x", -1, 0, 1e30, [], {}, [1, 2], {"otype": "Identity"}, "HashGrid", 3.5, 1 << 31, "Grid", 64, 128, 7, 33, 1e-9, [[1, 2, 3, 4]] * 4, [[1, 2], [3]]]

BASE_CONFIG = {
    "loss": {"otype": "Huber"},
    "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                                                           "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
    "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 19, "base_resolution": 16},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1},
    "dir_encoding": {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity", "n_bins": 4, "degree": 4}]},
    "rgb_network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}


def _mutate(rng, o, p_delete, p_replace):
    if isinstance(o, dict):
        for k in list(o.keys()):
            r = rng.random()
            if r < p_delete:
                del o[k]
            elif r < p_delete + p_replace:
                o[k] = copy.deepcopy(rng.choice(JUNK))
            else:
                _mutate(rng, o[k], p_delete, p_replace)
    elif isinstance(o, list):
        for i in range(len(o)):
            if rng.random() < p_replace:
                o[i] = copy.deepcopy(rng.choice(JUNK))
            else:
                _mutate(rng, o[i], p_delete, p_replace)


def test_factories_survive_mutated_configs(ngp):
    lib = ngp._lib.lib
    rng = random.Random(5)
    made = refused = 0
    for _ in range(200):
        j = copy.deepcopy(BASE_CONFIG)
        _mutate(rng, j, 0.05, 0.15)
        enc, net = json.dumps(j.get("encoding", {})).encode(), json.dumps(j.get("network", {})).encode()
        calls = ((lib.ngpb_create_network_with_input_encoding, (3, 4, enc, net)), (lib.ngpb_create_encoding, (3, enc, 1)),
                 (lib.ngpb_create_nerf_network, (3, 3, 0, 4, enc, json.dumps(j.get("dir_encoding", {})).encode(), net, json.dumps(j.get("rgb_network", {})).encode())),
                 (lib.ngpb_optimizer_create, (json.dumps(j.get("optimizer", {})).encode(),)))
        for fn, args in calls:
            h = ctypes.c_void_p()
            if fn(*args, ctypes.byref(h)) != 0:
                assert lib.ngpb_last_error().decode() != ""
                refused += 1
                continue
            made += 1
            if fn is lib.ngpb_optimizer_create:
                lib.ngpb_optimizer_free(h)
            else:
                lib.ngpb_module_n_params(h)
                assert lib.ngpb_module_hyperparams(h, None, 0) > 0
                lib.ngpb_module_free(h)
    assert made > 100 and refused > 100


def test_dataset_describer_survives_mutated_transforms(ngp, tmp_path):
    lib = ngp._lib.lib
    rng = random.Random(6)
    frames = []
    for i in range(6):
        name = "images/%04d.jpg" % (i + 1)
        os.makedirs(tmp_path / "images", exist_ok=True)
        (tmp_path / name).write_bytes(b"")
        frames.append({"file_path": name, "sharpness": 30.0 + i, "transform_matrix": [[1, 0, 0, 0.1 * i], [0, 1, 0, 0.2], [0, 0, 1, 0.3], [0, 0, 0, 1]]})
    base = {"camera_angle_x": 0.8, "camera_angle_y": 1.2, "fl_x": 1375.5, "fl_y": 1374.5, "k1": 0.03, "k2": -0.007, "p1": 1e-4, "p2": 2e-4, "cx": 554.5, "cy": 965.9,
            "w": 1080, "h": 1920, "aabb_scale": 4, "frames": frames}
    extra_keys = ["scale", "offset", "up", "aabb_scale", "render_aabb", "n_frames", "sharpness_discard_threshold", "k1", "fl_x", "w", "h", "cx", "normal_mts_args",
                  "rolling_shutter", "is_fisheye", "white_transparent", "fix_premult"]
    described = refused = 0
    for _ in range(300):
        j = copy.deepcopy(base)
        _mutate(rng, j, 0.004, 0.008)
        if rng.random() < 0.3:
            j[rng.choice(extra_keys)] = copy.deepcopy(rng.choice(JUNK))
        (tmp_path / "transforms.json").write_text(json.dumps(j))
        n = lib.ngpb_host_describe_dataset(str(tmp_path / "transforms.json").encode(), None, 0)
        if n:
            buf = ctypes.create_string_buffer(n)
            lib.ngpb_host_describe_dataset(str(tmp_path / "transforms.json").encode(), buf, n)
            json.loads(buf.value.decode())
            described += 1
        else:
            refused += 1
    assert described > 50 and refused > 20
