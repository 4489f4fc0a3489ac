"""CPU tests: the oracle against the integer invariants derivable from the reference's code
(the reference ships no golden vectors), and the published PCG32
known-answer vector."""
import json
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

L16F2 = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19,
         "base_resolution": 16, "per_level_scale": 1.3819}
FOX_L8F4 = {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 19,
            "base_resolution": 16, "per_level_scale": float(np.exp(np.log(2048 * 4 / 16) / 7))}
ALBERT_2D = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 24,
             "base_resolution": 16, "per_level_scale": float(np.exp(np.log(512 / 16) / 15))}


def test_level_tables_match_golden(oracle):
    with open(os.path.join(GOLDEN, "grid_levels.json")) as f:
        golden = json.load(f)
    for case in golden["cases"]:
        cfg = oracle.grid_cfg(case["n_dims"], case["encoding"])
        offs = oracle.grid_offsets(cfg)
        _, res = oracle.grid_levels(cfg)
        assert int(offs[-1]) == case["n_entries"], case["name"]
        assert int(offs[-1]) * case["encoding"]["n_features_per_level"] == case["n_params"]
        if "resolutions_head" in case:
            assert list(map(int, res[: len(case["resolutions_head"])])) == case["resolutions_head"]
        if "n_dense_levels" in case:
            dense = 0
            for l in range(cfg.n_levels):
                if int(res[l]) ** case["n_dims"] <= int(offs[l + 1] - offs[l]):
                    dense += 1
            assert dense == case["n_dense_levels"], case["name"]


def test_pcg32_published_vector(oracle):
    # pcg32-demo (pcg-random.org, pcg32_srandom(42, 54)): first six 32-bit outputs. Our
    # binding only exposes floats from stream 1, so check the float mapping on seed 1337
    # against an independent Python restatement instead, and the demo vector through it.
    def pcg32_py(initstate, initseq, n, advance=0):
        M = (1 << 64) - 1
        mult = 0x5851f42d4c957f2d
        state, inc = 0, ((initseq << 1) | 1) & M

        def step():
            nonlocal state
            old = state
            state = (old * mult + inc) & M
            xs = (((old >> 18) ^ old) >> 27) & 0xFFFFFFFF
            rot = old >> 59
            return ((xs >> rot) | (xs << ((-rot) & 31))) & 0xFFFFFFFF

        step()
        state = (state + initstate) & M
        step()
        for _ in range(advance):
            step()
        return [step() for _ in range(n)]

    assert pcg32_py(42, 54, 6) == [0xa15c02b7, 0x7b47f409, 0xba1d3330, 0x83d2f293, 0xbfa4784b, 0xcbed606e]
    ref = pcg32_py(1337, 1, 8, advance=1000)
    exp = (np.array([(r >> 9) | 0x3f800000 for r in ref], dtype=np.uint32).view(np.float32) - np.float32(1.0))
    got = oracle.pcg32_floats(1337, 1000, 8)
    assert np.array_equal(got, exp)


def test_hash_of_fixed_coordinates(oracle):
    """grid_index on hand-evaluated coordinates (tcnn common_device.h:651-655, 690-707)."""
    cfg = oracle.grid_cfg(3, L16F2)
    offs = oracle.grid_offsets(cfg)
    scales, res = oracle.grid_levels(cfg)
    x = np.array([[0.3, 0.6, 0.9], [0.0, 0.0, 0.0], [0.999, 0.5, 0.25]], dtype=np.float32)
    idx = oracle.grid_indices(cfg, x)
    for i in range(x.shape[0]):
        for level in (0, 3, 7, 15):
            size = int(offs[level + 1] - offs[level])
            pos = np.float32(scales[level]) * x[i] + np.float32(0.5)  # fmaf differs from mul+add only in rounding;
            pg = np.floor(pos.astype(np.float64)).astype(np.int64)   # checked away from cell borders below
            frac = pos - np.floor(pos)
            if np.any(frac < 1e-3) or np.any(frac > 1 - 1e-3):
                continue
            for k in range(8):
                c = [int(pg[d]) + ((k >> d) & 1) for d in range(3)]
                r = int(res[level])
                if r ** 3 <= size:
                    want = (c[0] + c[1] * r + c[2] * r * r) % size
                else:
                    want = ((c[0] * 1) ^ ((c[1] * 2654435761) & 0xFFFFFFFF) ^ ((c[2] * 805459861) & 0xFFFFFFFF)) % size
                assert int(idx[i, level, k]) == want, (i, level, k)


def test_forward_is_linear_in_params_and_partition_of_unity(oracle):
    """Constant table -> every level returns that constant (weights sum to 1 up to fp16 rounding)."""
    cfg = oracle.grid_cfg(3, FOX_L8F4)
    n_entries = int(oracle.grid_offsets(cfg)[-1])
    rng = np.random.default_rng(0)
    x = rng.random((256, 3), dtype=np.float32)
    params = np.full(n_entries * 4, 0.5, dtype=np.float16)
    out = oracle.grid_forward(cfg, x, params).astype(np.float32)
    assert np.all(np.abs(out - 0.5) <= 2e-3)


def test_backward_is_adjoint_of_forward(oracle):
    """<forward(x; P), dY> == <P, backward(x; dY)> up to fp16 rounding of each product."""
    enc = dict(L16F2, n_levels=4, log2_hashmap_size=12)
    cfg = oracle.grid_cfg(3, enc)
    n_entries = int(oracle.grid_offsets(cfg)[-1])
    rng = np.random.default_rng(1)
    x = rng.random((512, 3), dtype=np.float32)
    params = rng.standard_normal(n_entries * 2).astype(np.float16)
    dy = rng.standard_normal((512, 8)).astype(np.float16)
    y = oracle.grid_forward(cfg, x, params).astype(np.float64)
    g = oracle.grid_backward(cfg, x, dy)
    lhs = float((y * dy.astype(np.float64)).sum())
    rhs = float((params.astype(np.float64) * g).sum())
    assert abs(lhs - rhs) <= 2e-2 * max(1.0, abs(lhs))


def test_input_gradient_matches_finite_difference(oracle):
    enc = dict(L16F2, n_levels=3, interpolation="Smoothstep")
    cfg = oracle.grid_cfg(3, enc)
    n_entries = int(oracle.grid_offsets(cfg)[-1])
    rng = np.random.default_rng(2)
    x = (rng.random((64, 3)) * 0.8 + 0.1).astype(np.float32)
    params = (rng.standard_normal(n_entries * 2) * 0.5).astype(np.float16)
    dy = rng.standard_normal((64, 6)).astype(np.float16)
    g = oracle.grid_backward_input(cfg, x, params, dy)
    eps = np.float32(2e-3)
    for d in range(3):
        xp, xm = x.copy(), x.copy()
        xp[:, d] += eps
        xm[:, d] -= eps
        fp = oracle.grid_forward(cfg, xp, params).astype(np.float64)
        fm = oracle.grid_forward(cfg, xm, params).astype(np.float64)
        fd = ((fp - fm) * dy.astype(np.float64)).sum(axis=1) / (2 * float(eps))
        ok = np.abs(fd - g[:, d]) <= 0.15 * np.abs(fd) + 0.3  # fp16 forward limits the finite difference
        assert ok.mean() > 0.9


def test_random_uniform_mapping(oracle):
    """generate_random_uniform's element -> stream position map (tcnn random.h:39-56)."""
    n = 1000
    got = oracle.random_uniform(1337, n, -1e-4, 1e-4)
    n_threads = 128 * (((n + 3) // 4 + 127) // 128)
    for e in (0, 1, 255, 256, 511, 999):
        t, j = e % n_threads, e // n_threads
        u = oracle.pcg32_floats(1337, 4 * t + j, 1)[0]
        want = np.float32(np.float64(u) * np.float64(np.float32(1e-4) - np.float32(-1e-4)) + np.float64(np.float32(-1e-4)))
        assert abs(got[e] - want) <= 1e-11


def test_ld_random_val_is_a_scrambled_sobol_sequence(oracle):
    # Sobol dimension 0 is the radical inverse: unscrambled values of consecutive indices are k / 2^32 bit-reversed
    v = [oracle.olib.oracle_ld_random_val(i, 12345, 0) for i in range(256)]
    assert all(0.0 <= x < 1.0 for x in v)
    # an Owen-scrambled (0,1)-sequence stays stratified: 256 points hit each of 256 strata exactly once
    assert sorted(int(x * 256) for x in v) == list(range(256))
    v1 = [oracle.olib.oracle_ld_random_val(i, 999, 1) for i in range(64)]
    assert sorted(int(x * 64) for x in v1) == list(range(64))


def _morton_xyz(oracle, idx):
    return np.array([[oracle.olib.oracle_morton3d_invert(int(i) >> s) for s in (0, 1, 2)] for i in idx], dtype=np.int64)


def test_occupancy_oracle_properties(oracle):
    """The occupancy-maintenance restatements (mark_untrained, grid samples, splat, EMA) obey what
    the reference code implies: samples lie inside the cell they index, only cells above the
    threshold are chosen when one is found, cells behind every camera are marked -1, the splat is a
    maximum and the EMA never revives an untrained cell."""
    rng = np.random.default_rng(3)
    # one camera at (0.5, -1.5, 0.5) looking along +y at the unit cube (NGP convention: columns right, down, forward, origin)
    cam = np.array([[1, 0, 0, 0.5], [0, 0, 1, -1.5], [0, -1, 0, 0.5]], dtype=np.float32)
    im = oracle.nerf_image(np.zeros((30, 40, 4), dtype=np.uint8), cam, focal=60.0)
    n_casc = 2
    marked = oracle.nerf_mark_untrained(np.zeros(n_casc * 128 ** 3, dtype=np.float32), [im])
    assert set(np.unique(marked).tolist()) == {-1.0, 0.0}
    xyz = _morton_xyz(oracle, np.arange(0, 128 ** 3, 4099))
    centre = (xyz + 0.5) / 128.0
    vis0 = marked[np.arange(0, 128 ** 3, 4099)] == 0
    # cascade 0 = the unit cube, in front of the camera: a cell is visible iff it projects into the image
    d = centre - np.array([0.5, -1.5, 0.5])
    u, v = d[:, 0] / d[:, 1] * 60.0 / 40.0 + 0.5, -d[:, 2] / d[:, 1] * 60.0 / 30.0 + 0.5
    clearly_in = (u > 0.05) & (u < 0.95) & (v > 0.05) & (v < 0.95)
    clearly_out = (u < -0.05) | (u > 1.05) | (v < -0.05) | (v > 1.05)
    assert vis0[clearly_in].all() and not vis0[clearly_out].any() and clearly_in.any() and clearly_out.any()
    # cascade 1 spans [-0.5, 1.5]^3: its cells with y < -1.5 do not exist, but those behind the image plane direction do
    cells1 = 128 ** 3 + np.arange(0, 128 ** 3, 4099)
    c1 = (_morton_xyz(oracle, cells1 - 128 ** 3) + 0.5) / 128.0 * 2.0 - 0.5
    d1 = c1 - np.array([0.5, -1.5, 0.5])
    assert (d1[:, 1] > 0).all()   # everything in [-0.5, 1.5]^3 is in front of a camera at y = -1.5

    grid = np.zeros(n_casc * 128 ** 3, dtype=np.float32)
    grid[rng.integers(0, grid.size, size=grid.size // 3)] = 0.5
    state = oracle.Pcg32State(11)
    n = 20000
    pos, idx = oracle.nerf_generate_grid_samples(n, state, 3, [-0.5] * 3, [1.5] * 3, grid, n_casc, 0.01)
    level = idx // 128 ** 3
    assert set(np.unique(level).tolist()) == {0, 1}
    world = pos[:, :3] * 2.0 - 0.5                       # unwarp_position for aabb [-0.5, 1.5]^3
    cell = np.floor(((world - 0.5) / (2.0 ** level)[:, None] + 0.5) * 128.0).astype(np.int64)
    want = _morton_xyz(oracle, idx % 128 ** 3)
    assert (np.abs(cell - want) <= 0).mean() > 0.999      # fp rounding may put a sample on a cell's upper face
    assert (grid[idx] > 0.01).mean() > 0.95               # ten tries at p = 1/3 find an occupied cell 98 % of the time
    raw = rng.normal(0, 1, size=n).astype(np.float16)
    tmp = oracle.nerf_splat_grid_samples(idx, raw, np.zeros_like(grid), 3)
    best = np.zeros_like(grid)
    np.maximum.at(best, idx, np.exp(raw.astype(np.float32)) * np.float32(np.sqrt(3.0) / 1024.0))
    np.testing.assert_allclose(tmp, best, rtol=1e-5)
    g = grid.copy()
    g[::7] = -1.0
    out = oracle.nerf_ema_grid_samples(0.9, g, tmp)
    assert (out[::7] == -1.0).all()
    keep = g >= 0
    np.testing.assert_array_equal(out[keep], np.maximum(g[keep] * np.float32(0.9), tmp[keep]))
