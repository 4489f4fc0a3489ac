"""Parity of the NeRF driver kernels with the CPU oracle, called through the C ABI.

Bars (BASELINE.json north_star): per-ray sample counts, sample coordinates and occupancy bitfields
BIT-EXACT; loss gradients within fp16 rounding (they are stored as fp16) plus 1e-3 relative for the
device's __expf / powf against the host libm.
"""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _look_at(eye, target=(0.5, 0.5, 0.5), up=(0.0, 0.0, 1.0)):
    eye = np.asarray(eye, dtype=np.float64)
    f = np.asarray(target) - eye
    f /= np.linalg.norm(f)
    r = np.cross(f, up)
    r /= np.linalg.norm(r)
    d = np.cross(f, r)
    return np.stack([r, d, f, eye], axis=1).astype(np.float32)


def _scene(oracle, rng, n_images, res, lens=None, aabb_scale=1, with_mask=False):
    """Random RGBA8 images around the unit cube + a random-but-structured occupancy grid."""
    w, h = res
    images = []
    for i in range(n_images):
        px = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
        if with_mask and i == 0:
            px[: h // 2, :, :] = np.array([255, 0, 255, 0], dtype=np.uint8)  # 0x00FF00FF = masked away
        ang = 2 * np.pi * i / n_images
        eye = np.array([0.5 + 1.3 * np.cos(ang), 0.5 + 1.3 * np.sin(ang), 0.5 + 0.4 * np.sin(3 * ang)])
        images.append(oracle.nerf_image(px, _look_at(eye), focal=1.2 * w, principal=(0.5 + 0.01 * i, 0.5), lens=lens))
    # occupancy: blobs of density in every cascade
    max_cascade = int(np.log2(aabb_scale))
    grid = np.zeros((max_cascade + 1, 128 ** 3), dtype=np.float32)
    for c in range(max_cascade + 1):
        idx = rng.integers(0, 128 ** 3 // 512, size=600) * 512
        for s in idx:
            grid[c, s:s + 512] = rng.random(512).astype(np.float32) * 0.5  # contiguous Morton runs = 8x8x8 bricks
    return images, grid.reshape(-1), max_cascade


def _config(oracle, n_rays, n_images, aabb_scale, max_cascade, max_samples, batch):
    cfg = oracle.NerfConfig()
    half = 0.5 * aabb_scale
    for k in range(3):
        cfg.aabb_min[k] = 0.5 - half
        cfg.aabb_max[k] = 0.5 + half
        cfg.background_color[k] = 0.0
    cfg.n_rays, cfg.n_rays_total = n_rays, 0
    cfg.max_samples, cfg.max_samples_compacted = max_samples, batch
    cfg.n_training_images, cfg.max_mip = n_images, max_cascade
    cfg.cone_angle_constant = 0.0 if aabb_scale <= 1 else 1.0 / 256.0
    cfg.near_distance, cfg.loss_scale = 0.1, 128.0
    cfg.snap_to_pixel_centers, cfg.random_bg_color, cfg.linear_colors, cfg.color_space = 1, 1, 0, 0
    cfg.loss_type, cfg.rgb_activation, cfg.density_activation = 4, 2, 3  # Huber, Logistic, Exponential
    return cfg


class _Device:
    """Device copies of the scene and the buffers of ngpb_nerf_train_buffers_t."""

    def __init__(self, ngp, oracle, cfg, images, bitfield, mean_density):
        self.keep = []
        meta = (oracle.NerfImage * len(images))()
        for i, im in enumerate(images):
            ctypes.memmove(ctypes.byref(meta[i]), ctypes.byref(im), ctypes.sizeof(oracle.NerfImage))
            px = _dev(im._keep)
            self.keep.append(px)
            meta[i].pixels = px.data_ptr()
        self.meta = _dev(np.frombuffer(bytes(meta), dtype=np.uint8).copy())
        self.bitfield = _dev(bitfield)
        self.mean = _dev(np.array([mean_density], dtype=np.float32))
        n, ms, mc = cfg.n_rays, cfg.max_samples, cfg.max_samples_compacted
        z = lambda shape, dt: torch.zeros(shape, dtype=dt, device="cuda")
        self.counters = z(4, torch.int32)
        self.ray_indices = z(n, torch.int32)
        self.rays = z((n, 6), torch.float32)
        self.numsteps = z((n, 2), torch.int32)
        self.coords = z((ms, 7), torch.float32)
        self.rgbsigma = z((ms, 4), torch.float16)
        self.coords_compacted = z((mc, 7), torch.float32)
        self.dloss = z((mc, 4), torch.float16)
        self.loss = z(n, torch.float32)
        b = oracle.NerfBuffers()
        b.metadata, b.density_bitfield, b.mean_density = self.meta.data_ptr(), self.bitfield.data_ptr(), self.mean.data_ptr()
        b.counters, b.ray_indices, b.rays_unnormalized = self.counters.data_ptr(), self.ray_indices.data_ptr(), self.rays.data_ptr()
        b.numsteps, b.coords, b.rgbsigma = self.numsteps.data_ptr(), self.coords.data_ptr(), self.rgbsigma.data_ptr()
        b.coords_compacted, b.dloss_doutput, b.loss = self.coords_compacted.data_ptr(), self.dloss.data_ptr(), self.loss.data_ptr()
        self.compact_src = z(mc, torch.int32)
        b.compact_src = self.compact_src.data_ptr()
        self.buf = b


def _generate(ngp, oracle, cfg, dev, rng_state):
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    st = (ctypes.c_uint64 * 2)(int(rng_state.arr[0]), int(rng_state.arr[1]))
    ngp._lib.check(ngp.lib.ngpb_nerf_generate_training_samples(stream, ctypes.byref(cfg), ctypes.byref(dev.buf), st))
    torch.cuda.synchronize()
    numsteps_total, _, n_rays_kept, _ = dev.counters.cpu().numpy().view(np.uint32)
    return int(numsteps_total), int(n_rays_kept)


@pytest.mark.parametrize("aabb_scale,lens,n_rays", [(1, None, 4096), (4, None, 4096), (4, (0.05, -0.02, 0.001, -0.002), 2048), (16, None, 1024)])
def test_training_samples_bit_exact(ngp, oracle, aabb_scale, lens, n_rays):
    rng = np.random.default_rng(aabb_scale * 7 + n_rays)
    images, grid, max_cascade = _scene(oracle, rng, 5, (64, 48), lens=lens, aabb_scale=aabb_scale, with_mask=True)
    bitfield, mean = oracle.nerf_bitfield(max_cascade, grid[:128 ** 3] if max_cascade == 0 else grid)
    max_samples = 1 << 21
    cfg = _config(oracle, n_rays, len(images), aabb_scale, max_cascade, max_samples, 1 << 20)
    step_rng = oracle.Pcg32State(1337)
    want_n, want_rays, want_coords, want_total = oracle.nerf_generate_samples(cfg, images, bitfield, step_rng, max_samples)
    assert want_total > 1000 and want_total < max_samples
    assert (want_n == 0).any() and (want_n > 0).any()  # masked / empty rays exist and are dropped

    dev = _Device(ngp, oracle, cfg, images, bitfield, mean)
    total, kept = _generate(ngp, oracle, cfg, dev, step_rng)
    assert total == want_total
    assert kept == int((want_n > 0).sum())
    ray_indices = dev.ray_indices.cpu().numpy().view(np.uint32)[:kept]
    numsteps = dev.numsteps.cpu().numpy().view(np.uint32)[:kept]
    assert sorted(ray_indices.tolist()) == np.nonzero(want_n)[0].tolist()
    # per-ray sample counts: bit-exact
    assert np.array_equal(numsteps[:, 0], want_n[ray_indices])
    # sample ranges tile [0, total) without overlap
    order = np.argsort(numsteps[:, 1])
    assert numsteps[order[0], 1] == 0
    assert np.array_equal(numsteps[order, 1][1:], np.cumsum(numsteps[order, 0])[:-1])
    # coordinates and rays: bit-exact, ray by ray
    want_base = np.concatenate([[0], np.cumsum(want_n.astype(np.int64))])[:-1].astype(np.int64)
    coords = dev.coords.cpu().numpy()
    rays = dev.rays.cpu().numpy()
    for k in range(kept):
        i, n, b = int(ray_indices[k]), int(numsteps[k, 0]), int(numsteps[k, 1])
        got = coords[b:b + n].view(np.uint32)
        want = want_coords[want_base[i]:want_base[i] + n].view(np.uint32)
        assert np.array_equal(got, want), (k, i)
        assert np.array_equal(rays[k].view(np.uint32), want_rays[i].view(np.uint32))


def test_training_samples_bit_exact_at_the_bench_shape(ngp, oracle):
    """configs[2] shape: 1080 x 1920 portrait views with the fox dataset's OpenCV lens and focal length, aabb_scale 4, the
    exponential stepping of cone angle 1/256, a 4096-ray slice of a 2^18-sample step: per-ray counts and coordinates
    bit-exact against the oracle (the small-image cases above cannot see index arithmetic that only overflows or rounds
    differently at full resolution)."""
    rng = np.random.default_rng(2024)
    fox_lens = (0.0578421, -0.0805099, -0.000980296, 0.00015575)   # data/nerf/fox/transforms.json k1 k2 p1 p2
    images, grid, max_cascade = _scene(oracle, rng, 3, (1080, 1920), lens=fox_lens, aabb_scale=4)
    for im in images:   # fox intrinsics (fl_x 1375.52, fl_y 1374.49, cx 554.558 / 1080, cy 965.268 / 1920)
        im.focal_x, im.focal_y, im.principal_x, im.principal_y = 1375.52, 1374.49, 554.558 / 1080, 965.268 / 1920
    bitfield, mean = oracle.nerf_bitfield(max_cascade, grid)
    n_rays, max_samples = 4096, 1 << 22
    cfg = _config(oracle, n_rays, len(images), 4, max_cascade, max_samples, 1 << 18)
    step_rng = oracle.Pcg32State(1337)
    step_rng.advance(77 << 32)   # a later training step
    want_n, want_rays, want_coords, want_total = oracle.nerf_generate_samples(cfg, images, bitfield, step_rng, max_samples)
    assert want_total > 50000 and want_n.max() > 100
    dev = _Device(ngp, oracle, cfg, images, bitfield, mean)
    total, kept = _generate(ngp, oracle, cfg, dev, step_rng)
    assert total == want_total and kept == int((want_n > 0).sum())
    ray_indices = dev.ray_indices.cpu().numpy().view(np.uint32)[:kept]
    numsteps = dev.numsteps.cpu().numpy().view(np.uint32)[:kept]
    assert np.array_equal(numsteps[:, 0], want_n[ray_indices])
    want_base = np.concatenate([[0], np.cumsum(want_n.astype(np.int64))])[:-1].astype(np.int64)
    coords = dev.coords.cpu().numpy()
    for k in range(kept):
        i, n, b = int(ray_indices[k]), int(numsteps[k, 0]), int(numsteps[k, 1])
        assert np.array_equal(coords[b:b + n].view(np.uint32), want_coords[want_base[i]:want_base[i] + n].view(np.uint32)), (k, i)


def test_training_samples_overflow_drops_whole_rays(ngp, oracle):
    rng = np.random.default_rng(3)
    images, grid, max_cascade = _scene(oracle, rng, 3, (32, 32))
    bitfield, mean = oracle.nerf_bitfield(0, grid)
    cfg = _config(oracle, 4096, 3, 1, 0, 4096, 4096)  # far too small a sample buffer
    dev = _Device(ngp, oracle, cfg, images, bitfield, mean)
    total, kept = _generate(ngp, oracle, cfg, dev, oracle.Pcg32State(5))
    numsteps = dev.numsteps.cpu().numpy().view(np.uint32)[:kept]
    assert total > cfg.max_samples and kept > 0
    assert (numsteps[:, 0] + numsteps[:, 1] <= cfg.max_samples).all()


@pytest.mark.parametrize("aabb_scale,loss_type,rgb_act", [(1, 4, 2), (4, 0, 2), (1, 6, 3)])
def test_loss_and_gradients_match_oracle(ngp, oracle, aabb_scale, loss_type, rgb_act):
    rng = np.random.default_rng(11 + aabb_scale)
    images, grid, max_cascade = _scene(oracle, rng, 4, (48, 48), aabb_scale=aabb_scale)
    bitfield, mean = oracle.nerf_bitfield(max_cascade, grid)
    n_rays, max_samples, batch = 2048, 1 << 20, 1 << 20
    cfg = _config(oracle, n_rays, len(images), aabb_scale, max_cascade, max_samples, batch)
    cfg.loss_type, cfg.rgb_activation = loss_type, rgb_act
    step_rng = oracle.Pcg32State(99)
    dev = _Device(ngp, oracle, cfg, images, bitfield, mean)
    total, kept = _generate(ngp, oracle, cfg, dev, step_rng)
    # a synthetic network output: moderate densities so that some rays terminate early, some do not
    # Keyed by (ray index, sample index), not by buffer position: sample slots are handed out by an atomicAdd in
    # block-arrival order, so a positional fill would give every run a different ray -> density assignment.
    ray_indices = dev.ray_indices.cpu().numpy().view(np.uint32)[:kept].copy()
    numsteps_in = dev.numsteps.cpu().numpy().view(np.uint32)[:kept].copy()
    out = np.zeros((max_samples, 4), dtype=np.float16)
    for k in range(kept):
        n, b = int(numsteps_in[k, 0]), int(numsteps_in[k, 1])
        r = np.random.default_rng([11 + aabb_scale, int(ray_indices[k])])
        out[b:b + n, :3] = r.normal(0, 1.0, size=(n, 3)).astype(np.float16)
        out[b:b + n, 3] = r.normal(2.0, 2.5, size=n).astype(np.float16)
    dev.rgbsigma.copy_(_dev(out))
    rays = dev.rays.cpu().numpy()[:kept].copy()
    coords = dev.coords.cpu().numpy()[:total].copy()

    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    st = (ctypes.c_uint64 * 2)(int(step_rng.arr[0]), int(step_rng.arr[1]))
    ngp._lib.check(ngp.lib.ngpb_nerf_compute_loss_and_compact(stream, ctypes.byref(cfg), ctypes.byref(dev.buf), st))
    torch.cuda.synchronize()

    comp, loss, dloss = oracle.nerf_compute_loss(cfg, images, mean, ray_indices, numsteps_in, rays, coords, out[:total], step_rng)
    got_ns = dev.numsteps.cpu().numpy().view(np.uint32)[:kept]
    n_compacted = int(dev.counters.cpu().numpy().view(np.uint32)[1])
    # compacted per-ray counts: bit-exact (early termination happens at the same sample: the exponentials of the
    # compositing chain are the same fixed polynomial on both sides, nerf_device.cuh composite_alpha)
    assert np.array_equal(got_ns[:, 0], comp)
    assert n_compacted == int(comp.sum())
    assert (comp < numsteps_in[:, 0]).any()
    np.testing.assert_allclose(dev.loss.cpu().numpy()[:kept], loss, rtol=2e-3, atol=1e-9)
    got_d = dev.dloss.cpu().numpy().astype(np.float32)
    got_c = dev.coords_compacted.cpu().numpy()
    got_src = dev.compact_src.cpu().numpy().view(np.uint32)
    scale = np.abs(dloss).max()
    for k in range(kept):
        n, cb, b = int(comp[k]), int(got_ns[k, 1]), int(numsteps_in[k, 1])
        assert np.array_equal(got_c[cb:cb + n].view(np.uint32), coords[b:b + n].view(np.uint32))
        assert np.array_equal(got_src[cb:cb + n], np.arange(b, b + n)), k   # the row each compacted sample was copied from (feature-tile reuse)
        want = dloss[b:b + n]
        # the density gradient is a suffix sum that cancels to exactly 0 at a ray's last sample in the
        # serial order; the warp-parallel suffix leaves an fp32 rounding residue there (a few 1e-7 of
        # a loss-scaled gradient), hence the absolute terms: half an fp16 ulp at the gradient scale
        # (the gradients are stored as halves) plus 1e-6
        tol = 2e-3 * np.abs(want) + 2.0 ** -10 * np.abs(want) + 2.0 ** -11 * scale + 1e-6
        err = np.abs(got_d[cb:cb + n] - want)
        assert (err <= tol).all(), (k, n, float((err / tol).max()), np.argwhere(err > tol)[:4].tolist(), got_d[cb:cb + n][err > tol][:4].tolist(), want[err > tol][:4].tolist())
    # roll-over: the batch is padded by wrapping around, padded gradients scaled by n / batch
    assert n_compacted < batch
    i = n_compacted + 17
    assert np.array_equal(got_c[i].view(np.uint32), got_c[i % n_compacted].view(np.uint32))
    assert got_src[i] == got_src[i % n_compacted]
    np.testing.assert_allclose(got_d[i], got_d[i % n_compacted] * n_compacted / batch, rtol=2e-3, atol=1e-7)


@pytest.mark.parametrize("max_cascade", [0, 2])
def test_bitfield_bit_exact(ngp, oracle, max_cascade):
    rng = np.random.default_rng(max_cascade)
    grid = (rng.random((max_cascade + 1) * 128 ** 3, dtype=np.float32) - 0.3).astype(np.float32) * 0.1
    grid[rng.integers(0, grid.size, size=1000)] = -1.0  # untrainable cells
    want_bf, want_mean = oracle.nerf_bitfield(max_cascade, grid)
    assert want_mean > 0.01  # threshold is then exactly NERF_MIN_OPTICAL_THICKNESS on both sides
    g = _dev(grid)
    bf = torch.zeros(ngp.lib.ngpb_nerf_bitfield_bytes(), dtype=torch.uint8, device="cuda")
    mean = torch.zeros(1, dtype=torch.float32, device="cuda")
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    ngp._lib.check(ngp.lib.ngpb_nerf_update_bitfield(stream, max_cascade, _ptr(g), _ptr(bf), _ptr(mean)))
    torch.cuda.synchronize()
    assert np.array_equal(bf.cpu().numpy(), want_bf)
    assert abs(float(mean.item()) - want_mean) <= 1e-5 * want_mean
    # coarser cascades are supersets (max pooling)
    assert want_bf[128 ** 3 // 8 * 7:].any()


def _meta_on_device(oracle, images):
    meta = (oracle.NerfImage * len(images))()
    keep = []
    for i, im in enumerate(images):
        ctypes.memmove(ctypes.byref(meta[i]), ctypes.byref(im), ctypes.sizeof(oracle.NerfImage))
        px = _dev(im._keep)
        keep.append(px)
        meta[i].pixels = px.data_ptr()
    return _dev(np.frombuffer(bytes(meta), dtype=np.uint8).copy()), keep


@pytest.mark.parametrize("aabb_scale,lens", [(1, None), (4, None), (2, (0.05, -0.02, 0.001, -0.002))])
def test_mark_untrained_density_grid_bit_exact(ngp, oracle, aabb_scale, lens):
    """mark_untrained_density_grid (src/testbed_nerf.cu:74-147): which cells any camera sees."""
    rng = np.random.default_rng(5 + aabb_scale)
    images, grid, max_cascade = _scene(oracle, rng, 3, (40, 30), lens=lens, aabb_scale=aabb_scale)
    n = 128 ** 3 * (max_cascade + 1)
    start = rng.normal(0.0, 1.0, size=n).astype(np.float32)   # both signs: the "keep unless the flag flips" branch is exercised
    meta, keep = _meta_on_device(oracle, images)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for clear in (1, 0):
        g = _dev(start)
        ngp._lib.check(ngp.lib.ngpb_nerf_mark_untrained_density_grid(stream, n, _ptr(g), len(images), _ptr(meta), clear))
        torch.cuda.synchronize()
        want = oracle.nerf_mark_untrained(start, images, clear_visible_voxels=bool(clear))
        got = g.cpu().numpy()
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), (clear, int((got != want).sum()))
        if clear:
            frac = float((got == 0).mean())
            assert 0.02 < frac < 0.98, frac   # three narrow cameras see part of the volume, not all of it


@pytest.mark.parametrize("aabb_scale", [1, 4])
def test_grid_samples_splat_and_ema_match_oracle(ngp, oracle, aabb_scale):
    """generate_grid_samples_nerf_nonuniform (bit-exact positions and cell indices), then
    splat_grid_samples_nerf_max_nearest_neighbor and ema_grid_samples_nerf on the same inputs."""
    rng = np.random.default_rng(17 + aabb_scale)
    _, grid, max_cascade = _scene(oracle, rng, 1, (8, 8), aabb_scale=aabb_scale)
    grid[rng.integers(0, grid.size, size=5000)] = -1.0   # untrained cells stay untouched by the EMA
    n_casc = max_cascade + 1
    half = 0.5 * aabb_scale
    aabb_min, aabb_max = [0.5 - half] * 3, [0.5 + half] * 3
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    state = oracle.Pcg32State(4242)
    st = (ctypes.c_uint64 * 2)(int(state.arr[0]), int(state.arr[1]))
    aabb = (ctypes.c_float * 6)(*(aabb_min + aabb_max))
    g_dev = _dev(grid)
    n = 100_003
    for step, thresh in ((0, -0.01), (7, 0.01)):
        pos = torch.zeros((n, 4), dtype=torch.float32, device="cuda")
        idx = torch.zeros(n, dtype=torch.int32, device="cuda")
        ngp._lib.check(ngp.lib.ngpb_nerf_generate_grid_samples(stream, n, st, step, aabb, _ptr(g_dev), _ptr(pos), _ptr(idx), n_casc, thresh))
        torch.cuda.synchronize()
        want_pos, want_idx = oracle.nerf_generate_grid_samples(n, state, step, aabb_min, aabb_max, grid, n_casc, thresh)
        assert np.array_equal(idx.cpu().numpy().view(np.uint32), want_idx)
        assert np.array_equal(pos.cpu().numpy().view(np.uint32), want_pos.view(np.uint32))
        assert (want_pos[:, :3] >= 0).all() and (want_pos[:, :3] <= 1).all() and (want_pos[:, 3] == 0).all()
        if thresh > 0:   # the hash retries up to 10 times for an occupied cell
            assert (grid[want_idx] > thresh).mean() > 0.05
    # splat: many samples per cell, the maximum wins; exact for None / ReLU, exp within the device's expf
    raw = rng.normal(0.0, 2.0, size=n).astype(np.float16)
    for act, exact in ((0, True), (1, True), (3, False)):
        tmp = torch.zeros(grid.size, dtype=torch.float32, device="cuda")
        if act == 0:
            raw_used = np.abs(raw)   # the bit-pattern maximum is only an order for non-negative values
        else:
            raw_used = raw
        ngp._lib.check(ngp.lib.ngpb_nerf_splat_grid_samples(stream, n, _ptr(idx), _ptr(_dev(raw_used)), _ptr(tmp), act))
        torch.cuda.synchronize()
        want_tmp = oracle.nerf_splat_grid_samples(want_idx, raw_used, np.zeros(grid.size, dtype=np.float32), act)
        got_tmp = tmp.cpu().numpy()
        if exact:
            assert np.array_equal(got_tmp, want_tmp)
        else:
            np.testing.assert_allclose(got_tmp, want_tmp, rtol=2e-6, atol=0)
        assert (got_tmp > 0).sum() > 1000
        # EMA on the device's own splat
        g2 = _dev(grid)
        ngp._lib.check(ngp.lib.ngpb_nerf_ema_grid_samples(stream, grid.size, 0.95, _ptr(g2), _ptr(tmp)))
        torch.cuda.synchronize()
        want_g = oracle.nerf_ema_grid_samples(0.95, grid, got_tmp)
        assert np.array_equal(g2.cpu().numpy().view(np.uint32), want_g.view(np.uint32))
        assert (want_g[grid < 0] == -1.0).all()
