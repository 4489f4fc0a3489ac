"""Parity of the tcgen05 fused MLP / NerfNetwork kernels against the CPU oracle (through the C ABI).

Tolerances (stated by north_star as "within a stated fp16 tolerance"):
  * network outputs: |gpu - oracle_fp32acc| <= 2^-8 * max|oracle| + 2^-12. Both round activations
    to fp16 between layers; the tensor core sums in a different order, so single fp16 roundings
    may flip (1 ulp = 2^-10 relative) and propagate through up to 5 layers.
  * the oracle's fp16-accumulate variant (model of the reference's wmma accumulator) is checked
    to lie within 2^-6 * max|.| of the fp32-accumulate oracle, i.e. the reference's own numerical
    noise is larger than our distance from the exact result.
  * gradients: relative L2 error <= 1e-2 after fp32/64 re-accumulation.
"""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

NERF_POS = {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.5}
NERF_DIR = {"otype": "Composite", "nested": [{"n_dims_to_encode": 3, "otype": "SphericalHarmonics", "degree": 4}, {"otype": "Identity"}]}
NERF_DENSITY = {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 1}
NERF_RGB = {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}


def _rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30)


def _close_fp16(got, want, rel=2.0 ** -8, abs_=2.0 ** -12):
    got = np.asarray(got, dtype=np.float32)
    want = np.asarray(want, dtype=np.float32)
    tol = rel * np.abs(want).max() + abs_
    err = np.abs(got - want).max()
    assert err <= tol, (err, tol)


def _weights(rng, in_width, n_hidden):
    """Xavier-like random weights, as fp16."""
    parts = []
    for rows, cols in [(64, in_width)] + [(64, 64)] * (n_hidden - 1) + [(16, 64)]:
        s = np.sqrt(6.0 / (rows + cols))
        parts.append(rng.uniform(-s, s, size=rows * cols))
    return np.concatenate(parts).astype(np.float16)


@pytest.mark.parametrize("in_dims,n_hidden,n", [(32, 2, 256), (16, 1, 512), (64, 3, 256), (32, 2, 1 << 14)])
def test_network_forward_matches_oracle(ngp, oracle, in_dims, n_hidden, n):
    rng = np.random.default_rng(in_dims * 10 + n_hidden)
    net = ngp.Network(in_dims, 3, {"otype": "FullyFusedMLP", "n_neurons": 64, "n_hidden_layers": n_hidden, "activation": "ReLU"})
    assert net.n_params == oracle.mlp_n_params(in_dims, n_hidden)
    assert net.layer_sizes() == [(64, in_dims)] + [(64, 64)] * (n_hidden - 1) + [(16, 64)]
    w = _weights(rng, in_dims, n_hidden)
    x = rng.uniform(-1, 1, size=(n, in_dims)).astype(np.float32)
    net.set_params(torch.from_numpy(w).cuda())
    out = net.inference(torch.from_numpy(x).cuda()).cpu().numpy()
    want, _ = oracle.mlp_forward(w, x.astype(np.float16), n_hidden)
    _close_fp16(out, want)
    # the reference's own accumulator noise is larger than that distance
    want16, _ = oracle.mlp_forward(w, x.astype(np.float16), n_hidden, accumulate_fp16=True)
    _close_fp16(want16, want, rel=2.0 ** -6)


def test_network_forward_activation_none(ngp, oracle):
    rng = np.random.default_rng(5)
    net = ngp.Network(32, 16, {"otype": "FullyFusedMLP", "n_neurons": 64, "n_hidden_layers": 2, "activation": "None"})
    w = _weights(rng, 32, 2)
    x = rng.uniform(-1, 1, size=(256, 32)).astype(np.float32)
    net.set_params(torch.from_numpy(w).cuda())
    out = net.inference(torch.from_numpy(x).cuda()).cpu().numpy()
    want, _ = oracle.mlp_forward(w, x.astype(np.float16), 2, relu=False)
    _close_fp16(out, want)


@pytest.mark.parametrize("in_dims,n_hidden,n", [(32, 2, 256), (16, 1, 256), (64, 3, 512), (32, 2, 1 << 15)])
def test_network_backward_matches_oracle(ngp, oracle, in_dims, n_hidden, n):
    rng = np.random.default_rng(100 + in_dims + n_hidden)
    net = ngp.Network(in_dims, 3, {"otype": "FullyFusedMLP", "n_neurons": 64, "n_hidden_layers": n_hidden})
    w = _weights(rng, in_dims, n_hidden)
    x = rng.uniform(-1, 1, size=(n, in_dims)).astype(np.float32)
    dy = (rng.standard_normal((n, 16)) * 0.5).astype(np.float16)
    net.set_params(torch.from_numpy(w).cuda())
    xt = torch.from_numpy(x).cuda()
    out, ctx = net.forward(xt)
    _, grad = net.backward(ctx, xt, out, torch.from_numpy(dy).cuda())
    torch.cuda.synchronize()
    _, hidden = oracle.mlp_forward(w, x.astype(np.float16), n_hidden)
    _, dW = oracle.mlp_backward(w, x.astype(np.float16), hidden, dy, n_hidden)
    assert _rel(grad.float().cpu().numpy(), dW) < 1e-2
    # per-matrix check so that a wrong small matrix cannot hide behind a large one
    offs = np.cumsum([0, 64 * in_dims] + [64 * 64] * (n_hidden - 1) + [16 * 64])
    g = grad.float().cpu().numpy()
    for a, b in zip(offs[:-1], offs[1:]):
        assert _rel(g[a:b], dW[a:b]) < 1e-2, (a, b)


def test_network_with_grid_encoding_matches_oracle(ngp, oracle):
    """Image-style network (configs/image/base.json shape, smaller table): 2-D grid -> 32 -> 64 -> 64 -> 3."""
    rng = np.random.default_rng(7)
    enc = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.26}
    netc = {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}
    m = ngp.NetworkWithInputEncoding(2, 3, enc, netc)
    n = 1024
    e = ngp.Encoding(2, enc)
    _, scales, _ = e.level_table()
    cfg = oracle.grid_cfg(2, enc, scales=scales)
    n_grid = e.n_params
    n_mlp = oracle.mlp_n_params(32, 2)
    assert m.n_params == n_grid + n_mlp
    w = _weights(rng, 32, 2)
    table = (rng.standard_normal(n_grid) * 0.3).astype(np.float16)
    params = np.concatenate([w, table])  # network first, then encoding (network_with_input_encoding.h:115-122)
    x = rng.random((n, 2), dtype=np.float32)
    dy = (rng.standard_normal((n, 16)) * 0.5).astype(np.float16)
    m.set_params(torch.from_numpy(params).cuda())
    xt = torch.from_numpy(x).cuda()
    out, ctx = m.forward(xt)
    _, grad = m.backward(ctx, xt, out, torch.from_numpy(dy).cuda())
    torch.cuda.synchronize()

    feat = oracle.grid_forward(cfg, x, table)
    want, hidden = oracle.mlp_forward(w, feat, 2)
    _close_fp16(out.cpu().numpy(), want)
    dfeat, dW = oracle.mlp_backward(w, feat, hidden, dy, 2)
    gg = oracle.grid_backward(cfg, x, dfeat)
    g = grad.float().cpu().numpy()
    assert _rel(g[:n_mlp], dW) < 1e-2
    assert _rel(g[n_mlp:], gg) < 2e-2


def _nerf_setup(ngp, oracle, rng, n, pos=NERF_POS):
    m = ngp.NerfNetwork(pos, NERF_DIR, NERF_DENSITY, NERF_RGB)
    e = ngp.Encoding(3, pos)
    _, scales, _ = e.level_table()
    cfg = oracle.grid_cfg(3, pos, scales=scales)
    nd, nr = oracle.mlp_n_params(32, 1), oracle.mlp_n_params(32, 2)
    assert m.n_params == nd + nr + e.n_params
    assert m.n_input_dims == 7 and m.n_output_dims == 16 and m.n_output_dims_unpadded == 4
    params = np.concatenate([_weights(rng, 32, 1), _weights(rng, 32, 2), (rng.standard_normal(e.n_params) * 0.3).astype(np.float16)])
    coords = rng.random((n, 7), dtype=np.float32)
    d = rng.standard_normal((n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    coords[:, 4:7] = (d + 1) * 0.5  # warp_direction (nerf_device.cuh:283-285)
    m.set_params(torch.from_numpy(params).cuda())
    return m, cfg, params, coords


def test_nerf_network_inference_matches_oracle(ngp, oracle):
    rng = np.random.default_rng(11)
    n = 2048
    m, cfg, params, coords = _nerf_setup(ngp, oracle, rng, n)
    out = m.inference(torch.from_numpy(coords).cuda()).cpu().numpy()
    fwd = oracle.nerf_forward(cfg, params, coords, 1, 2)
    _close_fp16(out[:, :4], fwd["out"][:, :4])
    _close_fp16(out[:, 4:], fwd["out"][:, 4:])
    fwd16 = oracle.nerf_forward(cfg, params, coords, 1, 2, accumulate_fp16=True)
    _close_fp16(fwd16["out"][:, :4], fwd["out"][:, :4], rel=2.0 ** -6)


def test_nerf_network_backward_matches_oracle(ngp, oracle):
    rng = np.random.default_rng(12)
    n = 4096
    m, cfg, params, coords = _nerf_setup(ngp, oracle, rng, n)
    dy = np.zeros((n, 16), dtype=np.float16)
    dy[:, :4] = (rng.standard_normal((n, 4)) * 0.5).astype(np.float16)
    xt = torch.from_numpy(coords).cuda()
    out, ctx = m.forward(xt)
    _, grad = m.backward(ctx, xt, out, torch.from_numpy(dy).cuda())
    torch.cuda.synchronize()
    fwd = oracle.nerf_forward(cfg, params, coords, 1, 2)
    dWd, dWr, gg, _ = oracle.nerf_backward(cfg, fwd, coords, dy, 1, 2)
    g = grad.float().cpu().numpy()
    nd, nr = fwd["nd"], fwd["nr"]
    assert _rel(g[:nd], dWd) < 1e-2
    assert _rel(g[nd:nd + nr], dWr) < 1e-2
    assert _rel(g[nd + nr:], gg) < 2e-2


def test_nerf_network_backward_is_deterministic_in_mlp_weights(ngp, oracle):
    """dW leaves TMEM once per CTA and is summed in a fixed order: bit-identical across runs."""
    rng = np.random.default_rng(13)
    n = 1 << 15
    m, cfg, params, coords = _nerf_setup(ngp, oracle, rng, n)
    dy = np.zeros((n, 16), dtype=np.float16)
    dy[:, :4] = (rng.standard_normal((n, 4)) * 0.5).astype(np.float16)
    xt, dyt = torch.from_numpy(coords).cuda(), torch.from_numpy(dy).cuda()
    out, ctx = m.forward(xt)
    n_mlp = oracle.mlp_n_params(32, 1) + oracle.mlp_n_params(32, 2)
    g1 = m.backward(ctx, xt, out, dyt)[1][:n_mlp].clone()
    g2 = m.backward(ctx, xt, out, dyt)[1][:n_mlp].clone()
    assert torch.equal(g1, g2)


def test_initialize_params_matches_reference_order(ngp, oracle):
    """density MLP, rgb MLP, pos encoding (nerf_network.h:369-381); xavier per matrix on the host
    generator, then the grid from the same generator (grid.h:1059-1062)."""
    m = ngp.NerfNetwork(NERF_POS, NERF_DIR, NERF_DENSITY, NERF_RGB)
    m.initialize_params(seed=1337)
    got = m.params_fp32.cpu().numpy()
    rng = oracle.Pcg32State(1337)
    wd = oracle.mlp_init(rng, 32, 1)
    wr = oracle.mlp_init(rng, 32, 2)
    grid = oracle.random_uniform_from(rng, m.n_params - wd.size - wr.size, -1e-4, 1e-4)
    want = np.concatenate([wd, wr, grid])
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_batch_size_granularity_is_enforced(ngp):
    net = ngp.Network(32, 3, {"otype": "FullyFusedMLP", "n_neurons": 64, "n_hidden_layers": 2})
    net.set_params(torch.zeros(net.n_params, dtype=torch.float16, device="cuda"))
    with pytest.raises(ngp.NgpbError):
        net.inference(torch.zeros((100, 32), device="cuda"))


def test_unsupported_width_raises(ngp):
    with pytest.raises(ngp.NgpbError, match="n_neurons"):
        ngp.Network(32, 3, {"otype": "FullyFusedMLP", "n_neurons": 128, "n_hidden_layers": 2})
