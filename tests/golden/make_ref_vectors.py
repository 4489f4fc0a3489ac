"""Writes tests/golden/ref_host_vectors.npz: outputs of the REFERENCE'S OWN header functions
(oracle/_ref/libref_host.so, built by oracle/build_ref.sh from /root/reference) on the seeded inputs
of tests/ref_pin_cases.py. Run in the authoring container (the reference does not travel):

    bash oracle/build_ref.sh && python tests/golden/make_ref_vectors.py
"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_pin_cases as rp  # noqa: E402
import oracle_binding as oracle  # noqa: E402  (level tables of the grid configs: inputs of the index probe)


def load_ref():
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_host.so"))
    lib.ref_probe.restype = ctypes.c_int
    return lib


def probe(fn, name, rows, k_out):
    rows = np.ascontiguousarray(rows, dtype=np.uint32)
    out = np.zeros((rows.shape[0], k_out), dtype=np.uint32)
    rc = fn(name.encode(), ctypes.c_uint32(rows.shape[0]), rows.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32)), ctypes.c_uint32(rows.shape[1]),
            out.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32)), ctypes.c_uint32(k_out))
    if rc != 0:
        raise KeyError(name)
    return out


def grid_rows(n_dims, enc):
    """(rows [n*L][D+3], n, L): every point at every level with that level's scale, resolution and table size."""
    cfg = oracle.grid_cfg(n_dims, enc)
    scales, res = oracle.grid_levels(cfg)
    offs = oracle.grid_offsets(cfg).astype(np.int64)
    x = rp.grid_points(n_dims)
    n, L = len(x), cfg.n_levels
    rows = np.zeros((n, L, n_dims + 3), dtype=np.uint32)
    rows[:, :, :n_dims] = x.view(np.uint32)[:, None, :]
    rows[:, :, n_dims] = scales.view(np.uint32)[None, :]
    rows[:, :, n_dims + 1] = res[None, :]
    rows[:, :, n_dims + 2] = (offs[1:] - offs[:-1]).astype(np.uint32)[None, :]
    return rows.reshape(n * L, n_dims + 3), n, L


def main():
    ref = load_ref()
    out = {}
    for name, (rows, k_out) in rp.cases().items():
        out[name] = probe(ref.ref_probe, name, rows, k_out)
    for name, (n_dims, enc) in rp.GRID_CONFIGS.items():
        rows, n, L = grid_rows(n_dims, enc)
        got = probe(ref.ref_probe, "grid_index%d" % n_dims, rows, 2 ** n_dims + n_dims)
        out["grid/" + name] = got.reshape(n, L, -1)
    path = os.path.join(HERE, "ref_host_vectors.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
