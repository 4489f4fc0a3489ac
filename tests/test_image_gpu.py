"""Neural-image mode (src/testbed_image.cu train_image, configs/image/base.json)
against numpy restatements of stratify2_kernel (:60-74), eval_image_kernel_and_snap (:164-210) and
tcnn's l2 / relative-l2 losses (losses/l2.h:38-75, relative_l2.h:38-76), called through the Testbed
C ABI. Positions are bit-exact (same pcg32 stream, power-of-two arithmetic); targets and loss
gradients within float / fp16 rounding; a fit of a synthetic image must converge."""
import importlib

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

IMAGE_CONFIG = {
    "loss": {"otype": "RelativeL2"},
    "optimizer": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 24, "base_resolution": 16},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}


def _image(w, h, seed=0):
    g = np.random.default_rng(seed)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.zeros((h, w, 4), dtype=np.float32)
    img[..., 0] = 0.5 + 0.5 * np.sin(xs / w * 9.0) * np.cos(ys / h * 5.0)
    img[..., 1] = ((xs // 16 + ys // 16) % 2) * 0.8 + 0.1          # checkerboard: sharp edges
    img[..., 2] = np.clip(xs / w + 0.05 * g.standard_normal((h, w)), 0, 1)
    img[..., 3] = 1.0
    return img


def _linear_to_srgb(l):
    l = l.astype(np.float32)
    return np.where(l < 0.0031308, 12.92 * l, 1.055 * np.power(np.maximum(l, 1e-20), np.float32(0.41666)) - 0.055).astype(np.float32)


@pytest.mark.parametrize("loss,snap", [("L2", True), ("RelativeL2", False)])
def test_image_training_step_matches_numpy(ngp, oracle, loss, snap):
    w, h, n = 200, 120, 1 << 16
    img = _image(w, h)
    tb = ngp.Testbed()
    tb.set_image(img)
    cfg = dict(IMAGE_CONFIG)
    cfg["loss"] = {"otype": loss}
    tb.reload_network_from_json(cfg)
    tb._set("image_snap_to_pixel_centers", int(snap))
    # derived grid parameters (src/testbed.cu:3694-3713): N_min = 2^(24/2), b = exp(ln(max(res)/2 / N_min) / 15)
    enc = tb.network_config["encoding"]
    assert enc["base_resolution"] == 16
    assert abs(enc["per_level_scale"] - np.exp(np.log(100.0 / 16) / 15)) < 1e-6
    tb.train(n)
    torch.cuda.synchronize()
    pos = tb.device_buffer("image_positions", torch.float32).cpu().numpy().reshape(n, 2)
    tgt = tb.device_buffer("image_targets", torch.float32).cpu().numpy().reshape(n, 3)
    out = tb.device_buffer("image_output", torch.float16).cpu().numpy().reshape(n, 16)
    dl = tb.device_buffer("image_dloss", torch.float16).cpu().numpy().reshape(n, 16)

    # positions: generate_random_uniform(m_rng = pcg32(1337)) then stratify2_kernel
    u = oracle.random_uniform(1337, 2 * n, 0.0, 1.0).reshape(n, 2)
    size = np.float32(256)
    i = np.arange(n)
    want = np.stack([u[:, 0] / size + (i & 255).astype(np.float32) / size, u[:, 1] / size + (i >> 8).astype(np.float32) / size], axis=1).astype(np.float32)
    res = np.array([w, h], dtype=np.float32)
    if snap:
        pi = np.floor(want * res)
        want_pos = ((pi + np.float32(0.5)) / res).astype(np.float32)
        pi = np.clip(pi.astype(np.int64), 0, [w - 1, h - 1])
        want_tgt = _linear_to_srgb(img[pi[:, 1], pi[:, 0], :3])
    else:
        want_pos = want
        p = np.clip(want * res - np.float32(0.5), 0, res - np.float32(1.0 + 1e-4)).astype(np.float32)
        p0 = p.astype(np.int64)
        wgt = p - p0.astype(np.float32)
        idx = np.clip(p0, 0, [w - 2, h - 2])
        s = _linear_to_srgb(img[..., :3])
        a, b = s[idx[:, 1], idx[:, 0]], s[idx[:, 1], idx[:, 0] + 1]
        c, d = s[idx[:, 1] + 1, idx[:, 0]], s[idx[:, 1] + 1, idx[:, 0] + 1]
        wx, wy = wgt[:, :1], wgt[:, 1:]
        want_tgt = (1 - wx) * (1 - wy) * a + wx * (1 - wy) * b + (1 - wx) * wy * c + wx * wy * d
    assert np.array_equal(pos.view(np.uint32), want_pos.view(np.uint32))
    np.testing.assert_allclose(tgt, want_tgt, rtol=2e-5, atol=2e-6)
    # every stratum of the 256 x 256 lattice holds exactly one sample
    cells = (np.floor(want[:, 0] * 256).astype(np.int64) + 256 * np.floor(want[:, 1] * 256).astype(np.int64))
    assert len(np.unique(cells)) == n

    # loss gradient: loss_scale * dloss/dprediction / (n * 3), zero on the 13 padding columns
    pred = out[:, :3].astype(np.float32)
    diff = pred - tgt
    grad = 2 * diff / (pred * pred + np.float32(0.01)) if loss == "RelativeL2" else 2 * diff
    want_dl = (np.float32(128.0) * grad / np.float32(n * 3)).astype(np.float16).astype(np.float32)
    got_dl = dl[:, :3].astype(np.float32)
    assert (np.abs(got_dl - want_dl) <= np.abs(want_dl) * 2.0 ** -9 + 6e-8).all()
    assert (dl[:, 3:] == 0).all()
    val = diff * diff / (pred * pred + np.float32(0.01)) if loss == "RelativeL2" else diff * diff
    assert abs(tb.loss - float(val.sum() / (n * 3))) <= 1e-3 * tb.loss
    tb.close()


def test_image_fit_converges_and_renders(ngp):
    w, h = 256, 256
    img = _image(w, h, seed=3)
    tb = ngp.Testbed()
    tb.set_image(torch.from_numpy(img).cuda().half())   # half-precision image path
    tb.reload_network_from_json(IMAGE_CONFIG)
    first = None
    for i in range(400):
        tb.train(1 << 16)
        if i == 0:
            first = tb.loss
    assert tb.training_step == 400
    for _ in range(16):
        tb.train(1 << 16)
    assert tb.loss < 0.05 * first, (first, tb.loss)
    got = tb.render_image(w, h).cpu().numpy()
    assert got.shape == (h, w, 4) and (got[..., 3] == 1).all()
    # compare in sRGB, the space the network is trained in (linear_colors = false)
    mse = float(((_linear_to_srgb(np.clip(got[..., :3], 0, 1)) - _linear_to_srgb(img[..., :3])) ** 2).mean())
    psnr = -10 * np.log10(mse)
    assert psnr > 27.0, psnr
    with pytest.raises(ngp.NgpbError):
        tb.train(1000)  # batch size granularity
    tb.close()
