"""Seeded inputs of the function-grain pin between the oracle and the reference's own header functions
(oracle/ref_host.cu). One table, used by tests/golden/make_ref_vectors.py (which runs the reference
functions in the authoring container and commits their outputs) and by tests/test_ref_pin_cpu.py.
Rows are 32-bit words; floats travel by bit pattern."""
import numpy as np


def _f(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _u(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


def _cat(*cols):
    return np.ascontiguousarray(np.concatenate([c.reshape(len(c), -1) for c in cols], axis=1), dtype=np.uint32)


def cases():
    """{name: (inputs uint32 [n][k_in], k_out)}"""
    g = np.random.default_rng(20240)
    n = 4096
    out = {}
    # positions in and around the aabb_scale 1..16 boxes centred at 0.5
    pos = (g.random((n, 3), dtype=np.float32) - 0.5) * np.float32(2.0) ** g.integers(0, 5, size=(n, 1)).astype(np.float32) + np.float32(0.5)
    out["cascaded_grid_idx_at"] = (_cat(_f(pos), _u(g.integers(0, 5, size=n))), 1)
    out["mip_from_pos"] = (_cat(_f(pos), _u(g.integers(0, 8, size=n))), 1)
    dt = np.exp(g.uniform(np.log(1e-4), np.log(0.5), size=n)).astype(np.float32)
    out["mip_from_dt"] = (_cat(_f(dt), _f(pos), _u(g.integers(0, 8, size=n))), 1)
    d = g.standard_normal((n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    res = (np.float32(128.0) / np.float32(2.0) ** g.integers(0, 5, size=n).astype(np.float32)).astype(np.float32)
    out["distance_to_next_voxel"] = (_cat(_f(pos), _f(d), _f(res)), 1)
    half = np.float32(2.0) ** g.integers(-1, 3, size=(n, 1)).astype(np.float32)
    bmin, bmax = np.float32(0.5) - half * np.ones((n, 3), np.float32), np.float32(0.5) + half * np.ones((n, 3), np.float32)
    o = (np.float32(0.5) + g.standard_normal((n, 3)).astype(np.float32) * np.float32(2.0)).astype(np.float32)
    out["ray_intersect"] = (_cat(_f(bmin), _f(bmax), _f(o), _f(d)), 2)
    lens = (g.standard_normal((n, 4)) * np.array([0.1, 0.03, 0.005, 0.005])).astype(np.float32)
    uv = (g.random((n, 2), dtype=np.float32) - np.float32(0.5)) * np.float32(1.2)
    out["opencv_undistort"] = (_cat(_f(lens), _f(uv)), 2)
    # cameras: rotation-ish 3x3 (any matrix will do: the map is linear) + origin, fox-like intrinsics, every other row an OpenCV lens
    cam = np.concatenate([g.standard_normal((n, 9)).astype(np.float32), (np.float32(0.5) + g.standard_normal((n, 3)).astype(np.float32))], axis=1)
    res = np.stack([g.integers(64, 4000, size=n), g.integers(64, 4000, size=n)], axis=1)
    focal = (res[:, :1] * g.uniform(0.6, 1.6, size=(n, 2))).astype(np.float32)
    pp = (0.5 + g.standard_normal((n, 2)) * 0.03).astype(np.float32)
    mode = (np.arange(n) % 2).astype(np.uint32)
    out["uv_to_ray"] = (_cat(_f(g.random((n, 2), dtype=np.float32)), _u(res), _f(focal), _f(pp), _f(cam), _u(mode), _f(lens)), 6)
    x01 = np.concatenate([g.random(n - 8, dtype=np.float32), np.array([0, 1, 0.04045, 0.0031308, 0.5, 1e-6, 0.04, 0.0031], np.float32)])
    out["srgb_to_linear"] = (_cat(_f(x01)), 1)
    out["linear_to_srgb"] = (_cat(_f(x01)), 1)
    tp = g.random((n, 6), dtype=np.float32) * np.float32(1.5) - np.float32(0.25)
    tp[::7, 3:] = tp[::7, :3] + (g.standard_normal((len(tp[::7]), 3)) * 0.03).astype(np.float32)   # inside the Huber knee
    out["loss"] = (_cat(_f(tp), _u(np.arange(n) % 7)), 6)
    v = (g.standard_normal(n) * 6).astype(np.float32)
    out["network_to_rgb"] = (_cat(_f(v), _u(np.arange(n) % 4)), 2)
    out["network_to_density"] = (_cat(_f(v), _u(np.arange(n) % 4)), 2)
    n_rays = g.integers(1, 1 << 20, size=n)
    out["image_idx"] = (_cat(_u(g.integers(0, n_rays)), _u(n_rays), _u(g.integers(1, 4000, size=n))), 1)
    out["ld_random_val"] = (_cat(_u(g.integers(0, 1 << 32, size=n, dtype=np.uint64)), _u(g.integers(0, 1 << 32, size=n, dtype=np.uint64)), _u(np.arange(n) % 2)), 1)
    out["warp_dt"] = (_cat(_f(dt)), 2)
    t = np.exp(g.uniform(np.log(1e-3), np.log(40.0), size=n)).astype(np.float32)
    cone = np.where(np.arange(n) % 3 == 0, np.float32(0.0), np.float32(1.0 / 256.0)).astype(np.float32)
    steps = g.uniform(0, 2000, size=n).astype(np.float32)
    out["stepping"] = (_cat(_f(t), _f(cone), _f(steps)), 4)
    xyz = g.integers(0, 1024, size=(n, 3))
    out["morton3D"] = (_u(xyz), 1)
    out["morton3D_invert"] = (_u(g.integers(0, 1 << 30, size=(n, 1))), 1)
    # ---- answered on the oracle side by its own exports (grid / pcg32 / SH) -------------------------
    lv = np.arange(n) % 16
    pls = np.array([1.3819, 1.5157, 2.4380271434783936, 1.26, 2.0], np.float32)[g.integers(0, 5, size=n)]
    out["grid_level"] = (_cat(_u(lv), _f(pls), _u(np.array([16, 16, 16, 8, 32])[g.integers(0, 5, size=n)])), 2)
    out["sh4"] = (_cat(_f(g.random((n, 3), dtype=np.float32))), 16)
    seeds = g.integers(0, 1 << 63, size=n, dtype=np.uint64)
    out["pcg32"] = (_cat(_u(seeds & np.uint64(0xFFFFFFFF)), _u(seeds >> np.uint64(32)), _u(g.integers(0, 1 << 24, size=n))), 2)
    return out


GRID_CONFIGS = {
    # name: (n_dims, encoding json) - configs[0] (both per_level_scales), this fork's NeRF grid, the image grid, the large table
    "L16F2_T19_aabb1": (3, {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.3819}),
    "L16F2_T19_aabb4": (3, {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.5157}),
    "L8F4_T19": (3, {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 2.4380271434783936}),
    "L16F2_T22": (3, {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 22, "base_resolution": 16, "per_level_scale": 1.5157}),
    "image_L16F2_T24": (2, {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 24, "base_resolution": 16, "per_level_scale": 1.2599}),
}


def grid_points(n_dims, n=1024):
    g = np.random.default_rng(77 + n_dims)
    x = g.random((n, n_dims), dtype=np.float32)
    x[:4] = np.array([[0.0] * n_dims, [1.0] * n_dims, [0.5] * n_dims, [np.float32(1.0) - np.float32(2.0 ** -24)] * n_dims], dtype=np.float32)
    return x


# ---- the render march of explicit rays (oracle_march_rays / ref_march_rays) -------------------------
def _morton_compact(v):
    v = v & np.uint32(0x09249249)
    v = (v ^ (v >> np.uint32(2))) & np.uint32(0x030C30C3)
    v = (v ^ (v >> np.uint32(4))) & np.uint32(0x0300F00F)
    v = (v ^ (v >> np.uint32(8))) & np.uint32(0xFF0000FF)
    v = (v ^ (v >> np.uint32(16))) & np.uint32(0x000003FF)
    return v


def march_scene(aabb_scale):
    """(density grid float32 [(max_cascade+1) * 128^3] in Morton order, max_cascade, aabb): a few solid blobs
    and a thin shell, so that rays see long empty stretches, grazing hits and dense interiors."""
    max_cascade = int(np.log2(aabb_scale))
    g = np.random.default_rng(5 + aabb_scale)
    centers = np.float32(0.5) + (g.random((7, 3)).astype(np.float32) - np.float32(0.5)) * np.float32(0.7 * aabb_scale)
    radii = (0.04 + 0.1 * g.random(7)).astype(np.float32) * np.float32(aabb_scale ** 0.5)
    idx = np.arange(128 ** 3, dtype=np.uint32)
    cell = np.stack([_morton_compact(idx), _morton_compact(idx >> np.uint32(1)), _morton_compact(idx >> np.uint32(2))], axis=1).astype(np.float32)
    grid = np.zeros((max_cascade + 1, 128 ** 3), dtype=np.float32)
    for k in range(max_cascade + 1):
        pos = ((cell + np.float32(0.5)) / np.float32(128.0) - np.float32(0.5)) * np.float32(2.0 ** k) + np.float32(0.5)
        occ = np.zeros(len(pos), dtype=bool)
        for c, r in zip(centers, radii):
            occ |= ((pos - c) ** 2).sum(axis=1) < r * r
        shell = np.abs(np.sqrt(((pos - np.float32(0.5)) ** 2).sum(axis=1)) - np.float32(0.45 * aabb_scale)) < np.float32(0.01 * aabb_scale)
        grid[k] = np.where(occ | shell, np.float32(10.0), np.float32(0.0))
    half = np.float32(0.5 * aabb_scale)
    aabb = np.array([0.5 - half] * 3 + [0.5 + half] * 3, dtype=np.float32)
    return grid.reshape(-1), max_cascade, aabb


def march_rays(aabb_scale, n=3000):
    """rays float32 [n][6] (origin, normalised direction) from outside and inside the box, and their pixel ids."""
    g = np.random.default_rng(31 + aabb_scale)
    o = np.float32(0.5) + g.standard_normal((n, 3)).astype(np.float32) * np.float32(0.6 * aabb_scale)
    target = np.float32(0.5) + (g.random((n, 3)).astype(np.float32) - np.float32(0.5)) * np.float32(0.9 * aabb_scale)
    d = target - o
    d = (d / np.sqrt((d * d).sum(axis=1, keepdims=True))).astype(np.float32)
    # renormalise in float32 the way normalize() would leave them: close enough to unit length for both sides, which
    # are handed the very same bits
    pix = g.integers(0, 1920 * 1080, size=n).astype(np.uint32)
    return np.ascontiguousarray(np.concatenate([o, d], axis=1), dtype=np.float32), pix


MARCH_CASES = [(1, 0.0, 0), (2, 1.0 / 256.0, 0), (4, 1.0 / 256.0, 3)]   # (aabb_scale, cone_angle_constant, spp_index)
