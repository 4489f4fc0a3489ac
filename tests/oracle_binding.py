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


def set_num_threads(n):
    olib.oracle_set_num_threads(ctypes.c_int(int(n)))


# ---- fused MLP / NerfNetwork glue (oracle/mlp_oracle.cpp) ---------------------------------------
olib.oracle_mlp_n_params.restype = ctypes.c_size_t
olib.oracle_mlp_n_params.argtypes = [ctypes.c_uint32] * 4
MLP_WIDTH = 64
MLP_PADDED_OUT = 16


def mlp_n_params(in_width, n_hidden, width=MLP_WIDTH, padded_out=MLP_PADDED_OUT):
    return int(olib.oracle_mlp_n_params(in_width, width, n_hidden, padded_out))


def mlp_forward(weights_f16, x_f16, n_hidden, relu=True, accumulate_fp16=False, width=MLP_WIDTH, padded_out=MLP_PADDED_OUT):
    """Returns (out [n][padded_out] fp16, hidden [n_hidden][n][width] fp16)."""
    w = np.ascontiguousarray(weights_f16, dtype=np.float16)
    x = np.ascontiguousarray(x_f16, dtype=np.float16)
    n, in_width = x.shape
    assert w.size == mlp_n_params(in_width, n_hidden, width, padded_out)
    out = np.zeros((n, padded_out), dtype=np.float16)
    hidden = np.zeros((n_hidden, n, width), dtype=np.float16)
    u16 = ctypes.c_uint16
    olib.oracle_mlp_forward(ctypes.c_uint32(n), ctypes.c_uint32(in_width), ctypes.c_uint32(width), ctypes.c_uint32(n_hidden),
                            ctypes.c_uint32(padded_out), ctypes.c_int(int(relu)), _p(w.view(np.uint16), u16),
                            _p(x.view(np.uint16), u16), ctypes.c_uint32(in_width), _p(out.view(np.uint16), u16),
                            ctypes.c_uint32(padded_out), _p(hidden.view(np.uint16), u16), ctypes.c_int(int(accumulate_fp16)))
    return out, hidden


def mlp_backward(weights_f16, x_f16, hidden_f16, dL_dout_f16, n_hidden, relu=True, width=MLP_WIDTH, padded_out=MLP_PADDED_OUT):
    """Returns (dL_dx [n][in] fp16, dW [n_params] float64)."""
    w = np.ascontiguousarray(weights_f16, dtype=np.float16)
    x = np.ascontiguousarray(x_f16, dtype=np.float16)
    h = np.ascontiguousarray(hidden_f16, dtype=np.float16)
    dy = np.ascontiguousarray(dL_dout_f16, dtype=np.float16)
    n, in_width = x.shape
    dx = np.zeros((n, in_width), dtype=np.float16)
    dW = np.zeros(w.size, dtype=np.float64)
    u16 = ctypes.c_uint16
    olib.oracle_mlp_backward(ctypes.c_uint32(n), ctypes.c_uint32(in_width), ctypes.c_uint32(width), ctypes.c_uint32(n_hidden),
                             ctypes.c_uint32(padded_out), ctypes.c_int(int(relu)), _p(w.view(np.uint16), u16),
                             _p(x.view(np.uint16), u16), ctypes.c_uint32(in_width), _p(h.view(np.uint16), u16),
                             _p(dy.view(np.uint16), u16), ctypes.c_uint32(dy.shape[1]), _p(dx.view(np.uint16), u16),
                             ctypes.c_uint32(in_width), _p(dW, ctypes.c_double))
    return dx, dW


class Pcg32State:
    """pcg32(seed) as {state, inc}; passed to the initialisers which advance it."""

    def __init__(self, seed):
        self.arr = np.zeros(2, dtype=np.uint64)
        olib.oracle_pcg32_seed(ctypes.c_uint64(seed), _p(self.arr, ctypes.c_uint64))

    def advance(self, delta):
        olib.oracle_pcg32_advance(_p(self.arr, ctypes.c_uint64), ctypes.c_int64(delta))


def mlp_init(rng, in_width, n_hidden, scale=1.0, width=MLP_WIDTH, padded_out=MLP_PADDED_OUT):
    out = np.zeros(mlp_n_params(in_width, n_hidden, width, padded_out), dtype=np.float32)
    olib.oracle_mlp_init(_p(rng.arr, ctypes.c_uint64), ctypes.c_uint32(in_width), ctypes.c_uint32(width), ctypes.c_uint32(n_hidden),
                         ctypes.c_uint32(padded_out), ctypes.c_float(scale), _p(out, ctypes.c_float))
    return out


def random_uniform_from(rng, n, lower, upper):
    """generate_random_uniform continuing from a generator state (tcnn random.h:39-73)."""
    out = np.zeros(n, dtype=np.float32)
    olib.oracle_random_uniform_state(_p(rng.arr, ctypes.c_uint64), ctypes.c_size_t(n), _p(out, ctypes.c_float), ctypes.c_float(lower), ctypes.c_float(upper))
    return out


def sh_encode(dirs_f32, degree=4):
    d = np.ascontiguousarray(dirs_f32, dtype=np.float32)
    out = np.zeros((d.shape[0], 16), dtype=np.float16)
    olib.oracle_sh_encode(ctypes.c_uint32(d.shape[0]), ctypes.c_uint32(degree), _p(d, ctypes.c_float), ctypes.c_uint32(d.shape[1]),
                          _p(out.view(np.uint16), ctypes.c_uint16), ctypes.c_uint32(16))
    return out


def nerf_forward(grid_cfg_, params_f16, coords, n_hidden_density, n_hidden_rgb, accumulate_fp16=False, sh_degree=4, dir_offset=4):
    """NerfNetwork::forward (nerf_network.h:145-187) composed from the oracle pieces.
    Returns dict with out [n][16] (rgb in 0-2, density-MLP output 0 in channel 3) and the intermediates."""
    L, F = grid_cfg_.n_levels, grid_cfg_.n_features_per_level
    pos_w = ((L * F + 15) // 16) * 16
    nd = mlp_n_params(pos_w, n_hidden_density)
    nr = mlp_n_params(32, n_hidden_rgb)
    p = np.ascontiguousarray(params_f16, dtype=np.float16)
    wd, wr, wg = p[:nd], p[nd:nd + nr], p[nd + nr:]
    x0 = grid_forward(grid_cfg_, coords[:, :3], wg, out_stride=pos_w)
    yd, hd = mlp_forward(wd, x0, n_hidden_density, accumulate_fp16=accumulate_fp16)
    sh = sh_encode(coords[:, dir_offset:dir_offset + 3], sh_degree)
    x1 = np.concatenate([yd, sh], axis=1)
    yr, hr = mlp_forward(wr, x1, n_hidden_rgb, accumulate_fp16=accumulate_fp16)
    out = yr.copy()
    out[:, 3] = yd[:, 0]
    return dict(out=out, x0=x0, yd=yd, hd=hd, x1=x1, yr=yr, hr=hr, wd=wd, wr=wr, nd=nd, nr=nr)


def nerf_backward(grid_cfg_, fwd, coords, dL_dout_f16, n_hidden_density, n_hidden_rgb):
    """NerfNetwork::backward (nerf_network.h:189-268). dL_dout [n][>=4] fp16: rgb 0-2, sigma 3.
    Returns float64 gradients (mlp density, mlp rgb, grid)."""
    dy = np.ascontiguousarray(dL_dout_f16, dtype=np.float16)
    n = dy.shape[0]
    g_rgb = np.zeros((n, 16), dtype=np.float16)
    g_rgb[:, :3] = dy[:, :3]
    dx1, dWr = mlp_backward(fwd["wr"], fwd["x1"], fwd["hr"], g_rgb, n_hidden_rgb)
    g_den = dx1[:, :16].astype(np.float32)
    g_den[:, 0] += dy[:, 3].astype(np.float32)
    g_den = g_den.astype(np.float16)
    dx0, dWd = mlp_backward(fwd["wd"], fwd["x0"], fwd["hd"], g_den, n_hidden_density)
    gg = grid_backward(grid_cfg_, coords[:, :3], dx0)
    return dWd, dWr, gg, dx0


# ---- NeRF driver kernels (oracle/nerf_oracle.cpp) -------------------------------------------------
class NerfImage(ctypes.Structure):
    """oracle_nerf_image == ngpb_nerf_image_t (same layout; pixels host vs device)."""
    _fields_ = [("pixels", ctypes.c_void_p), ("image_data_type", ctypes.c_int), ("lens_mode", ctypes.c_int), ("lens_params", ctypes.c_float * 7),
                ("res_x", ctypes.c_int), ("res_y", ctypes.c_int), ("principal_x", ctypes.c_float), ("principal_y", ctypes.c_float),
                ("focal_x", ctypes.c_float), ("focal_y", ctypes.c_float), ("xform", ctypes.c_float * 12)]


class NerfConfig(ctypes.Structure):
    """oracle_nerf_config == ngpb_nerf_train_config_t."""
    _fields_ = [("aabb_min", ctypes.c_float * 3), ("aabb_max", ctypes.c_float * 3),
                ("n_rays", ctypes.c_uint32), ("n_rays_total", ctypes.c_uint32), ("max_samples", ctypes.c_uint32), ("max_samples_compacted", ctypes.c_uint32),
                ("n_training_images", ctypes.c_uint32), ("max_mip", ctypes.c_uint32),
                ("cone_angle_constant", ctypes.c_float), ("near_distance", ctypes.c_float), ("loss_scale", ctypes.c_float),
                ("snap_to_pixel_centers", ctypes.c_int), ("random_bg_color", ctypes.c_int), ("linear_colors", ctypes.c_int), ("color_space", ctypes.c_int),
                ("loss_type", ctypes.c_int), ("rgb_activation", ctypes.c_int), ("density_activation", ctypes.c_int), ("background_color", ctypes.c_float * 3),
                ("ray_begin", ctypes.c_uint32), ("n_rays_local", ctypes.c_uint32)]


class NerfBuffers(ctypes.Structure):
    """ngpb_nerf_train_buffers_t (device pointers)."""
    _fields_ = [(n, ctypes.c_void_p) for n in ("metadata", "density_bitfield", "mean_density", "counters", "ray_indices", "rays_unnormalized",
                                               "numsteps", "coords", "rgbsigma", "coords_compacted", "dloss_doutput", "loss", "compact_src")]


olib.oracle_nerf_generate_samples.restype = ctypes.c_uint64
olib.oracle_det_logf.restype = ctypes.c_float
olib.oracle_det_logf.argtypes = [ctypes.c_float]
olib.oracle_det_expf.restype = ctypes.c_float
olib.oracle_det_expf.argtypes = [ctypes.c_float]
olib.oracle_advance_n_steps.restype = ctypes.c_float
olib.oracle_advance_n_steps.argtypes = [ctypes.c_float, ctypes.c_float, ctypes.c_float]
olib.oracle_morton3d.restype = ctypes.c_uint32
olib.oracle_morton3d.argtypes = [ctypes.c_uint32] * 3
olib.oracle_morton3d_invert.restype = ctypes.c_uint32
olib.oracle_morton3d_invert.argtypes = [ctypes.c_uint32]


def nerf_image(pixels_np, xform_3x4, focal, principal=(0.5, 0.5), lens=None):
    """Host-side image record; keeps the pixel array alive. xform_3x4 row-major [3][4] NGP convention."""
    px = np.ascontiguousarray(pixels_np)
    kind = {np.dtype(np.uint8): 1, np.dtype(np.float16): 2, np.dtype(np.float32): 3}[px.dtype]
    im = NerfImage()
    im._keep = px
    im.pixels = px.ctypes.data
    im.image_data_type = kind
    im.lens_mode = 1 if lens is not None and any(float(v) != 0.0 for v in lens) else 0
    for k in range(7):
        im.lens_params[k] = float(lens[k]) if lens is not None and k < len(lens) else 0.0
    im.res_x, im.res_y = px.shape[1], px.shape[0]
    im.principal_x, im.principal_y = principal
    im.focal_x, im.focal_y = (focal, focal) if np.isscalar(focal) else focal
    x = np.asarray(xform_3x4, dtype=np.float32)
    for c in range(4):
        for r in range(3):
            im.xform[3 * c + r] = float(x[r, c])
    return im


def nerf_generate_samples(cfg, images, bitfield, rng, max_samples):
    """Returns (numsteps [n_rays], rays [n_rays][6], coords [total][7], total) in ray-index order."""
    arr = (NerfImage * len(images))(*images)
    numsteps = np.zeros(cfg.n_rays, dtype=np.uint32)
    rays = np.zeros((cfg.n_rays, 6), dtype=np.float32)
    coords = np.zeros((max_samples, 7), dtype=np.float32)
    bf = np.ascontiguousarray(bitfield, dtype=np.uint8)
    total = olib.oracle_nerf_generate_samples(ctypes.byref(cfg), arr, _p(bf, ctypes.c_uint8), _p(rng.arr, ctypes.c_uint64),
                                              _p(numsteps, ctypes.c_uint32), _p(rays, ctypes.c_float), _p(coords, ctypes.c_float), ctypes.c_uint64(max_samples))
    return numsteps, rays, coords[:min(total, max_samples)], int(total)


def nerf_compute_loss(cfg, images, mean_density, ray_indices, numsteps_in, rays, coords, rgbsigma_f16, rng):
    """Returns (compacted_numsteps [n], loss [n], dloss [n_samples][4] float32 indexed by input sample)."""
    arr = (NerfImage * len(images))(*images)
    n = len(ray_indices)
    ri = np.ascontiguousarray(ray_indices, dtype=np.uint32)
    ns = np.ascontiguousarray(numsteps_in, dtype=np.uint32)
    ry = np.ascontiguousarray(rays, dtype=np.float32)
    co = np.ascontiguousarray(coords, dtype=np.float32)
    out = np.ascontiguousarray(rgbsigma_f16, dtype=np.float16)
    comp = np.zeros(n, dtype=np.uint32)
    loss = np.zeros(n, dtype=np.float32)
    dloss = np.zeros((co.shape[0], 4), dtype=np.float32)
    olib.oracle_nerf_compute_loss(ctypes.byref(cfg), arr, ctypes.c_float(mean_density), ctypes.c_uint32(n), _p(ri, ctypes.c_uint32), _p(ns, ctypes.c_uint32),
                                  _p(ry, ctypes.c_float), _p(co, ctypes.c_float), _p(out.view(np.uint16), ctypes.c_uint16), _p(rng.arr, ctypes.c_uint64),
                                  _p(comp, ctypes.c_uint32), _p(loss, ctypes.c_float), _p(dloss, ctypes.c_float))
    return comp, loss, dloss


def nerf_bitfield(max_cascade, grid):
    g = np.ascontiguousarray(grid, dtype=np.float32)
    bf = np.zeros(128 ** 3 * 8 // 8, dtype=np.uint8)
    mean = ctypes.c_float()
    olib.oracle_nerf_bitfield(ctypes.c_uint32(max_cascade), _p(g, ctypes.c_float), _p(bf, ctypes.c_uint8), ctypes.byref(mean))
    return bf, mean.value


# ---- occupancy-grid maintenance ------------------------------------------------------------------
def nerf_mark_untrained(grid, images, clear_visible_voxels=True):
    """In-place mark_untrained_density_grid on a copy of `grid`; returns the copy."""
    arr = (NerfImage * len(images))(*images)
    g = np.array(grid, dtype=np.float32, copy=True)
    olib.oracle_nerf_mark_untrained(ctypes.c_uint32(g.size), _p(g, ctypes.c_float), ctypes.c_uint32(len(images)), arr, ctypes.c_int(int(clear_visible_voxels)))
    return g


def nerf_generate_grid_samples(n, rng, step, aabb_min, aabb_max, grid_in, n_cascades, thresh):
    """Returns (positions [n][4] float32, indices [n] uint32)."""
    g = np.ascontiguousarray(grid_in, dtype=np.float32)
    aabb = np.array(list(aabb_min) + list(aabb_max), dtype=np.float32)
    pos = np.zeros((n, 4), dtype=np.float32)
    idx = np.zeros(n, dtype=np.uint32)
    olib.oracle_nerf_generate_grid_samples(ctypes.c_uint32(n), _p(rng.arr, ctypes.c_uint64), ctypes.c_uint32(step), _p(aabb, ctypes.c_float), _p(g, ctypes.c_float),
                                           _p(pos, ctypes.c_float), _p(idx, ctypes.c_uint32), ctypes.c_uint32(n_cascades), ctypes.c_float(thresh))
    return pos, idx


def nerf_splat_grid_samples(indices, density_raw_f16, grid_tmp, density_activation):
    g = np.array(grid_tmp, dtype=np.float32, copy=True)
    idx = np.ascontiguousarray(indices, dtype=np.uint32)
    d = np.ascontiguousarray(density_raw_f16, dtype=np.float16)
    olib.oracle_nerf_splat_grid_samples(ctypes.c_uint32(len(idx)), _p(idx, ctypes.c_uint32), _p(d.view(np.uint16), ctypes.c_uint16), _p(g, ctypes.c_float), ctypes.c_int(density_activation))
    return g


def nerf_ema_grid_samples(decay, grid, grid_tmp):
    g = np.array(grid, dtype=np.float32, copy=True)
    t = np.ascontiguousarray(grid_tmp, dtype=np.float32)
    olib.oracle_nerf_ema_grid_samples(ctypes.c_uint32(g.size), ctypes.c_float(decay), _p(g, ctypes.c_float), _p(t, ctypes.c_float))
    return g


def render_buffer_epilogue(frames, exposure=0.0, background=(0.0, 0.0, 0.0, 1.0), color_space=0, output_color_space=0, curve=0):
    """frames float32 [spp][n][4] -> [n][4] (accumulate_kernel + tonemap_kernel)."""
    f = np.ascontiguousarray(frames, dtype=np.float32)
    spp, n = f.shape[0], f.shape[1]
    bg = np.array(background, dtype=np.float32)
    out = np.zeros((n, 4), dtype=np.float32)
    olib.oracle_render_buffer_epilogue(ctypes.c_uint32(n), ctypes.c_uint32(spp), _p(f, ctypes.c_float), ctypes.c_float(exposure), _p(bg, ctypes.c_float),
                                       ctypes.c_int(color_space), ctypes.c_int(output_color_space), ctypes.c_int(curve), _p(out, ctypes.c_float))
    return out


def overlay_image(image_rgba8, width, height, fov_axis=1, zoom=1.0, exposure=0.0, background=(0.0, 0.0, 0.0, 1.0), color_space=0, output_color_space=0, curve=0):
    """numpy restatement of overlay_image_kernel with alpha = 1 (src/render_buffer.cu:348-411) for a
    Byte image (read_rgba, common_device.cuh:773-790): nearest texel along fov_axis, sRGB8 -> linear
    premultiplied, background, exposure, curve, transfer (the last three through the oracle's
    tonemap restatement). image_rgba8 uint8 [ih][iw][4] -> float32 [height*width][4]."""
    img = np.asarray(image_rgba8, dtype=np.uint8)
    ih, iw = img.shape[:2]
    f32 = np.float32
    scale = f32((iw, ih)[fov_axis]) / f32((width, height)[fov_axis])
    x = np.arange(width, dtype=np.float32) + f32(0.5)
    y = np.arange(height, dtype=np.float32) + f32(0.5)
    fx = (x - f32(width) * f32(0.5)) / f32(zoom) + f32(0.5) * f32(width)
    fy = (y - f32(height) * f32(0.5)) / f32(zoom) + f32(0.5) * f32(height)
    u = (fx - f32(width) * f32(0.5)) * scale + f32(iw) * f32(0.5)
    v = (fy - f32(height) * f32(0.5)) * scale + f32(ih) * f32(0.5)
    sx, sy = np.floor(u).astype(np.int64), np.floor(v).astype(np.int64)
    inside = ((sy >= 0) & (sy < ih))[:, None] & ((sx >= 0) & (sx < iw))[None, :]
    tex = img[np.clip(sy, 0, ih - 1)[:, None], np.clip(sx, 0, iw - 1)[None, :]].astype(np.float32) * f32(1.0 / 255.0)
    s = tex[..., :3]
    lin = np.where(s <= f32(0.04045), s / f32(12.92), ((s + f32(0.055)) / f32(1.055)) ** f32(2.4)).astype(np.float32)
    a = tex[..., 3:4]
    marker = (img[..., 0] == 0xFF) & (img[..., 1] == 0) & (img[..., 2] == 0xFF) & (img[..., 3] == 0)
    val = np.concatenate([lin * a, a], axis=-1)
    val[marker[np.clip(sy, 0, ih - 1)[:, None], np.clip(sx, 0, iw - 1)[None, :]]] = -1.0
    val[~inside] = 0.0
    assert color_space == 0, "restated for the linear accumulation space (the default; scripts/run.py's protocol)"
    # background, exposure, curve and transfer are those of tonemap_kernel: one frame, spp = 1
    return render_buffer_epilogue(val.reshape(1, -1, 4), exposure, background, color_space, output_color_space, curve)


# ---- NeRF rendering (oracle_nerf_render_*) --------------------------------------------------------
class RenderConfig(ctypes.Structure):
    """oracle_render_config."""
    _fields_ = [("aabb_min", ctypes.c_float * 3), ("aabb_max", ctypes.c_float * 3), ("camera", ctypes.c_float * 12),
                ("width", ctypes.c_uint32), ("height", ctypes.c_uint32),
                ("focal_x", ctypes.c_float), ("focal_y", ctypes.c_float), ("screen_cx", ctypes.c_float), ("screen_cy", ctypes.c_float),
                ("spp_index", ctypes.c_uint32), ("pixel_begin", ctypes.c_uint32), ("pixel_end", ctypes.c_uint32), ("max_mip", ctypes.c_uint32),
                ("cone_angle_constant", ctypes.c_float), ("min_transmittance", ctypes.c_float),
                ("rgb_activation", ctypes.c_int), ("density_activation", ctypes.c_int), ("linear_colors", ctypes.c_int), ("snap_to_pixel_centers", ctypes.c_int),
                ("max_steps_per_ray", ctypes.c_uint32), ("lens_mode", ctypes.c_int), ("lens_params", ctypes.c_float * 7)]


olib.oracle_nerf_render_march.restype = ctypes.c_uint64
olib.oracle_ld_random_val.restype = ctypes.c_float
olib.oracle_ld_random_val.argtypes = [ctypes.c_uint32] * 3


def render_config(aabb_min, aabb_max, camera_3x4, width, height, focal, max_mip, cone_angle_constant, principal=(0.5, 0.5), spp_index=0,
                  pixel_range=None, min_transmittance=0.01, rgb_activation=2, density_activation=3, linear_colors=False, snap=False, max_steps_per_ray=10000, lens=None):
    c = RenderConfig()
    if lens is not None:   # OpenCV k1, k2, p1, p2
        c.lens_mode = 1
        for k, val in enumerate(lens):
            c.lens_params[k] = float(val)
    for k in range(3):
        c.aabb_min[k], c.aabb_max[k] = float(aabb_min[k]), float(aabb_max[k])
    x = np.asarray(camera_3x4, dtype=np.float32)
    for col in range(4):
        for r in range(3):
            c.camera[3 * col + r] = float(x[r, col])
    c.width, c.height = width, height
    c.focal_x, c.focal_y = (focal, focal) if np.isscalar(focal) else focal
    c.screen_cx, c.screen_cy = principal
    c.spp_index = spp_index
    c.pixel_begin, c.pixel_end = pixel_range if pixel_range is not None else (0, width * height)
    c.max_mip = max_mip
    c.cone_angle_constant = cone_angle_constant
    c.min_transmittance = min_transmittance
    c.rgb_activation, c.density_activation = rgb_activation, density_activation
    c.linear_colors, c.snap_to_pixel_centers = int(linear_colors), int(snap)
    c.max_steps_per_ray = max_steps_per_ray
    return c


def nerf_render_march(cfg, bitfield):
    """Returns (numsteps [n][2] = count, base; coords [total][7])."""
    n = cfg.pixel_end - cfg.pixel_begin
    bf = np.ascontiguousarray(bitfield, dtype=np.uint8)
    numsteps = np.zeros((n, 2), dtype=np.uint32)
    total = olib.oracle_nerf_render_march(ctypes.byref(cfg), _p(bf, ctypes.c_uint8), _p(numsteps, ctypes.c_uint32), None, ctypes.c_uint64(0))
    coords = np.zeros((max(int(total), 1), 7), dtype=np.float32)
    olib.oracle_nerf_render_march(ctypes.byref(cfg), _p(bf, ctypes.c_uint8), _p(numsteps, ctypes.c_uint32), _p(coords, ctypes.c_float), ctypes.c_uint64(int(total)))
    return numsteps, coords[:int(total)]


def nerf_render_composite(cfg, numsteps, coords, rgbsigma_f16):
    """Returns (rgba [n][4], depth [n], steps [n])."""
    n = cfg.pixel_end - cfg.pixel_begin
    ns = np.ascontiguousarray(numsteps, dtype=np.uint32)
    co = np.ascontiguousarray(coords, dtype=np.float32)
    out = np.ascontiguousarray(rgbsigma_f16, dtype=np.float16)
    rgba = np.zeros((n, 4), dtype=np.float32)
    depth = np.zeros(n, dtype=np.float32)
    steps = np.zeros(n, dtype=np.uint32)
    olib.oracle_nerf_render_composite(ctypes.byref(cfg), _p(ns, ctypes.c_uint32), _p(co, ctypes.c_float), _p(out.view(np.uint16), ctypes.c_uint16),
                                      _p(rgba, ctypes.c_float), _p(depth, ctypes.c_float), _p(steps, ctypes.c_uint32))
    return rgba, depth, steps
