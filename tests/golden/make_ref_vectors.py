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


def march(fn, bitfield, rays, pix, spp_index, aabb, cone, max_mip, max_steps=4096):
    n = len(rays)
    counts, first, last = np.zeros(n, np.uint32), np.zeros(n, np.float32), np.zeros(n, np.float32)
    P = lambda a, t: a.ctypes.data_as(ctypes.POINTER(t))
    bf = np.ascontiguousarray(bitfield, dtype=np.uint8)
    fn.restype = None
    fn(ctypes.c_uint32(n), P(rays, ctypes.c_float), P(pix, ctypes.c_uint32), ctypes.c_uint32(spp_index), P(aabb, ctypes.c_float), ctypes.c_float(cone),
       ctypes.c_uint32(max_mip), P(bf, ctypes.c_uint8), ctypes.c_uint32(max_steps), P(counts, ctypes.c_uint32), P(first, ctypes.c_float), P(last, ctypes.c_float))
    return counts, first, last


def march_inputs(aabb_scale):
    grid, max_cascade, aabb = rp.march_scene(aabb_scale)
    bitfield, _ = oracle.nerf_bitfield(max_cascade, grid)
    rays, pix = rp.march_rays(aabb_scale)
    return bitfield, max_cascade, aabb, rays, pix


def main():
    ref = load_ref()
    out = {}
    for aabb_scale, cone, spp in rp.MARCH_CASES:
        bitfield, max_cascade, aabb, rays, pix = march_inputs(aabb_scale)
        counts, first, last = march(ref.ref_march_rays, bitfield, rays, pix, spp, aabb, cone, max_cascade)
        out["march/%d/counts" % aabb_scale], out["march/%d/first" % aabb_scale], out["march/%d/last" % aabb_scale] = counts, first, last
        out["march/%d/bitfield_crc" % aabb_scale] = np.array([int(np.frombuffer(bitfield.tobytes(), dtype=np.uint8).astype(np.uint64).sum())], dtype=np.uint64)
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
