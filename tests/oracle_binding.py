"""ctypes binding of oracle/liboracle.so for tests, smoke() and bench.py's CPU baseline.
TEST INFRASTRUCTURE: never imported by the product package."""
import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_LIB = os.path.join(ROOT, "oracle", "liboracle.so")
olib = ctypes.CDLL(ORACLE_LIB)

GRID_TYPES = {"hash": 0, "dense": 1, "tiled": 2, "tile": 2}
HASH_TYPES = {"prime": 0, "coherentprime": 1, "reversedprime": 2}
INTERP_TYPES = {"nearest": 0, "linear": 1, "smoothstep": 2}


class GridCfg(ctypes.Structure):
    _fields_ = [
        ("n_dims", ctypes.c_uint32), ("n_features_per_level", ctypes.c_uint32), ("n_levels", ctypes.c_uint32),
        ("log2_hashmap_size", ctypes.c_uint32), ("base_resolution", ctypes.c_uint32), ("per_level_scale", ctypes.c_float),
        ("grid_type", ctypes.c_int), ("hash_type", ctypes.c_int), ("interpolation", ctypes.c_int),
        ("scale_override", ctypes.POINTER(ctypes.c_float)),
    ]


def _p(a, t):
    return a.ctypes.data_as(ctypes.POINTER(t))


olib.oracle_grid_offsets.restype = ctypes.c_uint32
olib.oracle_float_to_half.restype = ctypes.c_uint16
olib.oracle_float_to_half.argtypes = [ctypes.c_float]
olib.oracle_half_to_float.restype = ctypes.c_float
olib.oracle_half_to_float.argtypes = [ctypes.c_uint16]
olib.oracle_num_threads.restype = ctypes.c_int


def grid_cfg(n_dims, enc, scales=None):
    """Builds the oracle's config from a tcnn encoding JSON dict (defaults: grid.h:1143-1176)."""
    otype = enc.get("otype", "Grid").lower()
    default_type = "tiled" if otype == "tiledgrid" else ("dense" if otype == "densegrid" else "hash")
    gtype = GRID_TYPES[enc.get("type", default_type).lower()]
    F = enc.get("n_features_per_level", 2)
    L = enc["n_features"] // F if "n_features" in enc else enc.get("n_levels", 16)
    base = enc.get("base_resolution", 16)
    if "per_level_scale" in enc:
        pls = np.float32(enc["per_level_scale"])
    elif gtype == 1:
        pls = np.exp(np.log(np.float32(256.0) / np.float32(base)) / np.float32(L - 1)).astype(np.float32)
    else:
        pls = np.float32(2.0)
    cfg = GridCfg(n_dims, F, L, enc.get("log2_hashmap_size", 19), base, float(pls), gtype,
                  HASH_TYPES[enc.get("hash", "CoherentPrime").lower()], INTERP_TYPES[enc.get("interpolation", "Linear").lower()], None)
    if scales is not None:
        arr = np.ascontiguousarray(scales, dtype=np.float32)
        cfg._keep = arr
        cfg.scale_override = _p(arr, ctypes.c_float)
    return cfg


def grid_offsets(cfg):
    offs = np.zeros(cfg.n_levels + 1, dtype=np.uint32)
    olib.oracle_grid_offsets(ctypes.byref(cfg), _p(offs, ctypes.c_uint32))
    return offs


def grid_levels(cfg):
    s = np.zeros(cfg.n_levels, dtype=np.float32)
    r = np.zeros(cfg.n_levels, dtype=np.uint32)
    olib.oracle_grid_levels(ctypes.byref(cfg), _p(s, ctypes.c_float), _p(r, ctypes.c_uint32))
    return s, r


def grid_indices(cfg, x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, stride = x.shape
    idx = np.zeros((n, cfg.n_levels, 2 ** cfg.n_dims), dtype=np.uint32)
    olib.oracle_grid_indices(ctypes.byref(cfg), ctypes.c_uint32(n), _p(x, ctypes.c_float), ctypes.c_uint32(stride), _p(idx, ctypes.c_uint32))
    return idx


def grid_forward(cfg, x, params_f16, out_stride=None, max_level=1.0):
    x = np.ascontiguousarray(x, dtype=np.float32)
    params = np.ascontiguousarray(params_f16, dtype=np.float16)
    n, stride = x.shape
    width = cfg.n_levels * cfg.n_features_per_level
    out_stride = out_stride or width
    out = np.zeros((n, out_stride), dtype=np.float16)
    olib.oracle_grid_forward(ctypes.byref(cfg), ctypes.c_uint32(n), _p(x, ctypes.c_float), ctypes.c_uint32(stride),
                             _p(params.view(np.uint16), ctypes.c_uint16), _p(out.view(np.uint16), ctypes.c_uint16),
                             ctypes.c_uint32(out_stride), ctypes.c_float(max_level))
    return out


def grid_backward(cfg, x, dL_dy_f16, max_level=1.0):
    x = np.ascontiguousarray(x, dtype=np.float32)
    dy = np.ascontiguousarray(dL_dy_f16, dtype=np.float16)
    n, stride = x.shape
    n_entries = int(grid_offsets(cfg)[-1])
    grad = np.zeros(n_entries * cfg.n_features_per_level, dtype=np.float64)
    olib.oracle_grid_backward(ctypes.byref(cfg), ctypes.c_uint32(n), _p(x, ctypes.c_float), ctypes.c_uint32(stride),
                              _p(dy.view(np.uint16), ctypes.c_uint16), ctypes.c_uint32(dy.shape[1]),
                              _p(grad, ctypes.c_double), ctypes.c_float(max_level))
    return grad


def grid_backward_input(cfg, x, params_f16, dL_dy_f16, max_level=1.0):
    x = np.ascontiguousarray(x, dtype=np.float32)
    params = np.ascontiguousarray(params_f16, dtype=np.float16)
    dy = np.ascontiguousarray(dL_dy_f16, dtype=np.float16)
    n, stride = x.shape
    out = np.zeros((n, cfg.n_dims), dtype=np.float32)
    olib.oracle_grid_backward_input(ctypes.byref(cfg), ctypes.c_uint32(n), _p(x, ctypes.c_float), ctypes.c_uint32(stride),
                                    _p(params.view(np.uint16), ctypes.c_uint16), _p(dy.view(np.uint16), ctypes.c_uint16),
                                    ctypes.c_uint32(dy.shape[1]), _p(out, ctypes.c_float), ctypes.c_float(max_level))
    return out


def pcg32_floats(seed, advance, n):
    out = np.zeros(n, dtype=np.float32)
    olib.oracle_pcg32_floats(ctypes.c_uint64(seed), ctypes.c_uint64(advance), ctypes.c_uint32(n), _p(out, ctypes.c_float))
    return out


def random_uniform(seed, n, lower, upper):
    out = np.zeros(n, dtype=np.float32)
    olib.oracle_random_uniform(ctypes.c_uint64(seed), ctypes.c_size_t(n), _p(out, ctypes.c_float), ctypes.c_float(lower), ctypes.c_float(upper))
    return out


def num_threads():
    return int(olib.oracle_num_threads())
