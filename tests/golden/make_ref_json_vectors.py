"""Writes tests/golden/ref_json_vectors.json.gz: what the REFERENCE'S OWN JSON library (nlohmann json vendored under
tiny-cuda-nn/dependencies/json, compiled by oracle/build_ref.sh into oracle/_ref/libref_json.so) produces for
  * json::to_msgpack of documents that exercise every encoding choice a snapshot meets (integer widths, float32-when-exact,
    string / array / map / bin size classes, key order), and of a snapshot-shaped document;
  * parent.merge_patch(child) of network configs (merge_parent_network_config, src/testbed.cu:86-97).
Run here (needs /root/reference): python tests/golden/make_ref_json_vectors.py"""
import ctypes
import json
import os

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_json.so"))
lib.ref_json_to_msgpack.restype = ctypes.c_size_t
lib.ref_json_to_msgpack.argtypes = [ctypes.c_char_p, ctypes.c_void_p, ctypes.c_size_t]
lib.ref_json_merge_patch.restype = ctypes.c_size_t
lib.ref_json_merge_patch.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_size_t]


def to_msgpack(text):
    n = lib.ref_json_to_msgpack(text.encode(), None, 0)
    buf = ctypes.create_string_buffer(n)
    lib.ref_json_to_msgpack(text.encode(), buf, n)
    return buf.raw[:n]


def merge(parent, child):
    n = lib.ref_json_merge_patch(parent.encode(), child.encode(), None, 0)
    buf = ctypes.create_string_buffer(n)
    lib.ref_json_merge_patch(parent.encode(), child.encode(), buf, n)
    return buf.value.decode()


docs = {
    "integers": json.dumps([0, 1, 127, 128, 255, 256, 65535, 65536, 4294967295, 4294967296, 9007199254740992, -1, -32, -33, -128, -129, -32768, -32769, -2147483648, -2147483649]),
    "floats": "[0.5, 0.25, 1.0, 0.1, 0.95, 1e-15, 1e-8, 0.33, 3.4e38, 1e300, -2.5, 0.009999999776482582, 1.5157166]",
    "strings": json.dumps(["", "a", "x" * 31, "y" * 32, "z" * 255, "w" * 256, "v" * 65536]),
    "containers": json.dumps({"a15": list(range(15)), "a16": list(range(16)), "m15": {"k%02d" % i: i for i in range(15)}, "m16": {"k%02d" % i: i for i in range(16)},
                              "nested": {"z": None, "b": True, "a": False, "list": [[1, 2], [3.5, "s"], {}], "empty": []}}),
    "binaries": json.dumps({"b0": "bin:", "b3": "bin:000102", "b255": "bin:" + "ab" * 255, "b256": "bin:" + "cd" * 256, "b70000": "bin:" + "ef" * 70000}),
    "key_order": '{"zeta": 1, "alpha": 2, "Beta": 3, "_x": 4, "10": 5, "9": 6}',
    "snapshot_shaped": json.dumps({
        "encoding": {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.5157166},
        "optimizer": {"otype": "Ema", "decay": 0.95, "nested": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                                                              "nested": {"otype": "Adam", "learning_rate": 0.01, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}}},
        "snapshot": {"version": 1, "mode": "nerf", "n_params": 12865536, "params_type": "__half", "params_binary": "bin:" + "3c00" * 64, "density_grid_size": 128,
                     "density_grid_binary": "bin:" + "0000" * 32, "training_step": 5000, "loss": 0.0005289230612106621, "bounding_radius": 1.0,
                     "aabb": {"min": [-1.5, -1.5, -1.5], "max": [2.5, 2.5, 2.5]}, "up_dir": [0.0, 1.0, 0.0], "exposure": 0.0, "background_color": [0.0, 0.0, 0.0, 1.0],
                     "camera": {"matrix": [[1.0, 0.0, 0.0, 0.5], [0.0, -1.0, 0.0, 0.5], [0.0, 0.0, -1.0, 0.5]], "fov_axis": 1, "relative_focal_length": [1.0, 1.0],
                                "screen_center": [0.5, 0.5], "zoom": 1.0, "scale": 1.5, "aperture_size": 0.0, "autofocus": False},
                     "nerf": {"aabb_scale": 4, "rgb": {"rays_per_batch": 4096, "measured_batch_size": 267968, "measured_batch_size_before_compaction": 310234},
                              "dataset": {"n_images": 1, "paths": ["images/0001.jpg"], "scale": 0.33, "offset": [0.5, 0.5, 0.5], "from_mitsuba": False, "is_hdr": False,
                                          "metadata": [{"focal_length": [1375.52, 1374.49], "principal_point": [0.5128, 0.5016], "resolution": [1080, 1920], "lens": None,
                                                        "rolling_shutter": [0.0, 0.0, 0.0, 0.0]}]}}}}),
}
merges = {
    "override_and_delete": ('{"loss": {"otype": "Huber"}, "encoding": {"otype": "HashGrid", "n_levels": 16, "log2_hashmap_size": 19}, "network": {"n_neurons": 64, "n_hidden_layers": 1}}',
                            '{"loss": null, "encoding": {"log2_hashmap_size": 21}, "network": {"n_hidden_layers": 2, "activation": "ReLU"}}'),
    "replace_non_object": ('{"optimizer": {"otype": "Adam", "learning_rate": 0.01}, "envmap": [1, 2, 3]}', '{"optimizer": "none", "envmap": {"resolution": [4, 4]}, "new": {"a": {"b": null, "c": 1}}}'),
    "with_comments": ('{ // parent\n "a": {"x": 1, "y": 2} /* c */ }', '{"a": {"y": null, "z": [1, {"k": null}]}}'),
}
out = {"_source": "nlohmann json vendored by the reference (tiny-cuda-nn/dependencies/json/json.hpp) through oracle/ref_json.cc; strings 'bin:<hex>' stand for json::binary_t",
       "to_msgpack": {k: {"json": v, "msgpack_hex": to_msgpack(v).hex()} for k, v in docs.items()},
       "merge_patch": {k: {"parent": p, "child": c, "merged": merge(p, c)} for k, (p, c) in merges.items()}}
import gzip
with gzip.GzipFile(os.path.join(HERE, "ref_json_vectors.json.gz"), "wb", mtime=0) as f:
    f.write(json.dumps(out).encode())
print({k: len(v["msgpack_hex"]) // 2 for k, v in out["to_msgpack"].items()}, os.path.getsize(os.path.join(HERE, "ref_json_vectors.json.gz")))
