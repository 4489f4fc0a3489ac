"""CPU tests of the boundary: the library loads, exports every symbol include/ngp_b200.h declares,
parses the tcnn JSON schema with the reference's defaults and error texts, and fails loudly
(no CPU fallback) when a compute entry point is called without a device."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "ngp_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ngpb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(ngp):
    lib = ctypes.CDLL(ngp.LIB_PATH)
    declared = _declared_symbols()
    assert len(declared) >= 25
    missing = [s for s in declared if not hasattr(lib, s)]
    assert missing == []
    assert ngp._lib.MISSING_SYMBOLS == []
    assert sorted(ngp._lib.SIGNATURES.keys()) == declared


def test_constants(ngp):
    assert ngp.batch_size_granularity() == 256   # tcnn common.h:235
    assert ngp.default_loss_scale() == 128.0     # tcnn common.h:232
    assert ngp.lib.ngpb_version() >= 100


def test_grid_config_defaults_and_table(ngp, oracle):
    enc = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19,
           "base_resolution": 16, "per_level_scale": 1.3819}
    e = ngp.Encoding(3, enc)
    assert e.n_input_dims == 3 and e.n_output_dims == 32
    assert e.n_params == 12196240
    offs, scales, res = e.level_table()
    cfg = oracle.grid_cfg(3, enc, scales=scales if ngp.has_device() else None)
    assert list(oracle.grid_offsets(cfg)) == offs
    o_scales, o_res = oracle.grid_levels(cfg)
    assert list(o_res) == res
    hp = e.hyperparams()
    assert hp["otype"] == "Grid" and hp["type"] == "Hash" and hp["hash"] == "CoherentPrime"
    assert hp["interpolation"] == "Linear" and hp["n_levels"] == 16 and hp["log2_hashmap_size"] == 19
    # defaults (grid.h:1145-1168): L=16, F=2, T=2^19, N_min=16, b=2
    d = ngp.Encoding(3, {"otype": "HashGrid"})
    hp = d.hyperparams()
    assert (hp["n_levels"], hp["n_features_per_level"], hp["base_resolution"], hp["log2_hashmap_size"]) == (16, 2, 16, 19)
    assert hp["per_level_scale"] == 2.0
    dense = ngp.Encoding(2, {"otype": "DenseGrid", "n_levels": 4})
    assert dense.hyperparams()["type"] == "Dense"


@pytest.mark.parametrize("enc,msg", [
    ({"otype": "HashGrid", "n_features_per_level": 3}, "n_features_per_level must be 1, 2, 4, or 8"),
    ({"otype": "HashGrid", "hash": "Nope"}, "Invalid hash type: Nope"),
    ({"otype": "HashGrid", "type": "Sparse"}, "Invalid grid type: Sparse"),
    ({"otype": "HashGrid", "interpolation": "Cubic"}, "Invalid interpolation type: Cubic"),
    ({"otype": "HashGrid", "n_features": 32, "n_levels": 16}, "may not specify n_features and n_levels simultaneously"),
    ({"otype": "Blob"}, "Invalid encoding type: Blob"),
])
def test_reference_error_texts(ngp, enc, msg):
    with pytest.raises(ngp.NgpbError) as ei:
        ngp.Encoding(3, enc)
    assert msg in str(ei.value)


def test_json_with_comments_and_bad_json(ngp):
    h = ctypes.c_void_p()
    txt = b'{ // comment\n "otype": "HashGrid", /* block */ "n_levels": 4 }'
    assert ngp.lib.ngpb_create_encoding(3, txt, 1, ctypes.byref(h)) == 0
    assert ngp.lib.ngpb_module_n_output_dims(h) == 8
    ngp.lib.ngpb_module_free(h)
    assert ngp.lib.ngpb_create_encoding(3, b'{"otype": ', 1, ctypes.byref(h)) == 1
    assert b"json parse error" in ngp.lib.ngpb_last_error()


def test_no_cpu_fallback(ngp):
    if ngp.has_device():
        pytest.skip("device present")
    e = ngp.Encoding(3, {"otype": "HashGrid", "n_levels": 2})
    x = np.zeros((256, 3), dtype=np.float32)
    out = np.zeros((256, 4), dtype=np.float16)
    params = np.zeros(e.n_params, dtype=np.float16)
    st = ngp.lib.ngpb_module_inference(e._h, None, 256, x.ctypes.data, out.ctypes.data, params.ctypes.data)
    assert st == 1
    assert b"no CUDA device" in ngp.lib.ngpb_last_error()


def test_product_does_not_reference_oracle():
    """The product path must never import, link or call anything under oracle/."""
    pkg = os.path.join(ROOT, "instant-ngp-rendering_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")) and f != "build.py":
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in text and "oracle_binding" not in text and "oracle/" not in text, f
