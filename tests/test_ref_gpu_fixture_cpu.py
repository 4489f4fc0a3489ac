"""The CPU oracle against the REFERENCE'S OWN GPU KERNELS: tests/golden/ref_gpu_vectors.npz holds what tiny-cuda-nn's
kernel_grid / kernel_grid_backward, kernel_mlp_fused (+ backward), ngp::NerfNetwork and adam_step / ema_step produced on a
B200 (baseline/dump_ref_gpu.py over the seeded inputs of tests/ref_gpu_cases.py). tests/test_ref_gpu_pin.py holds the CUDA
path to that file on the GPU; this file holds the ORACLE to it without one - the pin the oracle's parity claim rests on.

Bars (measured values in brackets):
  * grid forward: the oracle's fp16 hfma2 emulation is bit-identical to the reference kernel on >= 99 % of the features,
    the rest within 2 fp16 ulp [100 % identical on the three grids];
  * grid backward: the reference sums fp16 atomics in arbitrary order, the oracle sums in float64: rel-L2 <= 5e-3
    [1.6e-4 .. 4.0e-4], same set of touched entries;
  * MLP forward: against the oracle's fp16-accumulate variant (the reference's wmma accumulator) |d| <= 2^-7 max|.|
    [2.4e-4 .. 1.4e-3], against the fp32-accumulate variant <= 2^-6 max|.|;
  * MLP / NerfNetwork weight gradients: rel-L2 <= 4e-2 per matrix; NerfNetwork grid gradient rel-L2 <= 3e-2;
  * Adam + EMA after 24 steps (restated here in float32 numpy, expression for expression as adam.h:72-118 / ema.h:63-78):
    fp32 master weights rtol 1e-5, fp16 weights within 1 ulp, EMA >= 99.9 % within 1 ulp.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_gpu_cases as rc  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "ref_gpu_vectors.npz")


@pytest.fixture(scope="module")
def ref():
    return dict(np.load(GOLDEN))


def _device_scales(name):
    """Per-level scales as the GPU's exp2f gives them (tests/golden/device_level_scales.json, written on a B200 by
    tests/golden/make_device_scales.py): the reference's kernel computes them on the device, and the host's exp2f differs in
    the last bit on some levels."""
    import json
    with open(os.path.join(ROOT, "tests", "golden", "device_level_scales.json")) as f:
        return json.load(f)[name]


def _ulps16(a, b):
    def key(x):
        u = np.ascontiguousarray(x, dtype=np.float16).view(np.uint16).astype(np.int32)
        return np.where(u & 0x8000, -(u & 0x7FFF), u & 0x7FFF)
    return np.abs(key(a) - key(b))


def _rel(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def _dense(n, idx, val):
    d = np.zeros(n, dtype=np.float64)
    d[idx] = val.astype(np.float64)
    return d


@pytest.mark.parametrize("name", sorted(rc.GRID_CASES))
def test_oracle_grid_against_reference_kernels(oracle, ref, name):
    n_dims, enc, x, params, dy = rc.grid_case(name)
    cfg = oracle.grid_cfg(n_dims, enc, scales=_device_scales(name))
    got = oracle.grid_forward(cfg, x, params)
    u = _ulps16(got, ref["grid/%s/y" % name])
    print("oracle grid fwd", name, "identical", float((u == 0).mean()), "max ulp", int(u.max()))
    assert (u == 0).mean() >= 0.99 and u.max() <= 2
    g = oracle.grid_backward(cfg, x, dy)
    gw = _dense(len(params), ref["grid/%s/dparams_idx" % name], ref["grid/%s/dparams_val" % name])
    print("oracle grid bwd", name, "rel", _rel(g, gw))
    assert ((g != 0) != (gw != 0)).mean() < 1e-4
    assert _rel(g, gw) <= 5e-3


@pytest.mark.parametrize("name", sorted(rc.MLP_CASES))
def test_oracle_mlp_against_reference_kernels(oracle, ref, name):
    n_in, n_out, cfg, x, w, dy = rc.mlp_case(name)
    n_hidden = cfg["n_hidden_layers"]
    want = ref["mlp/%s/y" % name].astype(np.float32)[:, :n_out]
    y16, h16 = oracle.mlp_forward(w, x.astype(np.float16), n_hidden, accumulate_fp16=True)
    y32, h32 = oracle.mlp_forward(w, x.astype(np.float16), n_hidden)
    e16 = np.abs(y16.astype(np.float32)[:, :n_out] - want).max() / np.abs(want).max()
    e32 = np.abs(y32.astype(np.float32)[:, :n_out] - want).max() / np.abs(want).max()
    print("oracle mlp fwd", name, "fp16-acc", e16, "fp32-acc", e32)
    assert e16 <= 2.0 ** -7 and e32 <= 2.0 ** -6
    _, dW = oracle.mlp_backward(w, x.astype(np.float16), h32, dy, n_hidden)
    gw = ref["mlp/%s/dparams" % name].astype(np.float64)
    offs = np.cumsum([0, 64 * n_in] + [64 * 64] * (n_hidden - 1) + [16 * 64])
    for a, b in zip(offs[:-1], offs[1:]):
        print("oracle mlp bwd", name, (int(a), int(b)), _rel(dW[a:b], gw[a:b]))
        assert _rel(dW[a:b], gw[a:b]) <= 4e-2


def test_oracle_nerf_network_against_reference_kernels(oracle, ref):
    cfg, x, params, dy = rc.nerfnet_case()
    gcfg = oracle.grid_cfg(3, cfg["encoding"], scales=_device_scales("nerfnet"))
    fwd = oracle.nerf_forward(gcfg, params, x, 1, 2)
    got = fwd["out"].astype(np.float32)[:, :4]
    for key in ("nerfnet/y", "nerfnet/y_inference"):
        want = ref[key].astype(np.float32)
        err = np.abs(got - want).max(axis=0) / np.abs(want).max(axis=0)
        print("oracle", key, err.tolist())
        assert (err <= 2.0 ** -6).all()
    want_density = ref["nerfnet/density"].astype(np.float32).reshape(len(x), -1)[:, 0]
    assert np.abs(fwd["yd"].astype(np.float32)[:, 0] - want_density).max() <= 2.0 ** -6 * np.abs(want_density).max()
    dWd, dWr, gg, _ = oracle.nerf_backward(gcfg, fwd, x, dy, 1, 2)
    gw = ref["nerfnet/dparams_mlp"].astype(np.float64)
    nd = rc.mlp_n_params(32, 1)
    print("oracle nerfnet dW density", _rel(dWd, gw[:nd]), "rgb", _rel(dWr, gw[nd:]))
    assert _rel(dWd, gw[:nd]) <= 2e-2 and _rel(dWr, gw[nd:]) <= 2e-2
    n_mlp = nd + rc.mlp_n_params(32, 2)
    gw_grid = _dense(len(params) - n_mlp, ref["nerfnet/dparams_grid_idx"], ref["nerfnet/dparams_grid_val"])
    print("oracle nerfnet grid grad", _rel(gg, gw_grid))
    assert _rel(gg, gw_grid) <= 3e-2


def test_adam_ema_restatement_against_reference_kernels(ref):
    """adam_step (tcnn optimizers/adam.h:72-118) inside ExponentialDecay (exponential_decay.h:60-71) inside EMA
    (ema.h:63-78, 102-103) in float32 numpy, on the case the reference ran."""
    cfg, w, grads = rc.adam_case()
    f = np.float32
    n, n_matrix = cfg["n"], sum(r * c for r, c in cfg["layers"])
    adam = cfg["optimizer"]["nested"]["nested"]
    beta1, beta2, eps, lr0, l2 = f(adam["beta1"]), f(adam["beta2"]), f(adam["epsilon"]), f(adam["learning_rate"]), f(adam["l2_reg"])
    decay = f(cfg["optimizer"]["decay"])
    w32 = w.astype(np.float32).copy()
    w16 = w32.astype(np.float16)
    m1, m2 = np.zeros(n, f), np.zeros(n, f)
    steps = np.zeros(n, np.uint32)
    ema = np.zeros(n, np.float16)
    is_matrix = np.arange(n) < n_matrix
    for s in range(cfg["steps"]):
        g = grads[s].astype(np.float32) / f(cfg["loss_scale"])
        active = is_matrix | (g != 0)
        g = np.where(is_matrix, g + l2 * w32, g).astype(f)
        m1 = np.where(active, beta1 * m1 + (f(1) - beta1) * g, m1).astype(f)
        m2 = np.where(active, beta2 * m2 + (f(1) - beta2) * (g * g), m2).astype(f)
        steps = np.where(active, steps + 1, steps).astype(np.uint32)
        t = steps.astype(np.float32)
        with np.errstate(divide="ignore", invalid="ignore"):   # entries that have never stepped (t = 0) divide 0 by 0; they are not `active`
            lr = (lr0 * np.sqrt(f(1) - np.power(beta2, t, dtype=f)) / (f(1) - np.power(beta1, t, dtype=f))).astype(f)
            eff = (lr / (np.sqrt(m2) + eps)).astype(f)
        new_w = (w32 - eff * m1).astype(f)
        w32 = np.where(active, new_w, w32).astype(f)
        w16 = np.where(active, w32.astype(np.float16), w16)
        step = s + 1
        debias_old, debias_new = f(1) - f(float(decay) ** (step - 1)), f(1) / (f(1) - f(float(decay) ** step))
        ema = ((ema.astype(f) * decay * debias_old + w16.astype(f) * (f(1) - decay)) * debias_new).astype(f).astype(np.float16)
    np.testing.assert_allclose(w32, ref["adam/w32"], rtol=1e-5, atol=1e-7)
    u = _ulps16(w16, ref["adam/w16"])
    assert u.max() <= 1 and (u == 0).mean() > 0.99
    u = _ulps16(ema, ref["adam/ema"])
    print("restated ema identical", float((u == 0).mean()), "max ulp", int(u.max()))
    assert (u <= 1).mean() >= 0.999 and u.max() <= 8
