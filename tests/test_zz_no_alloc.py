"""ngpb_module_inference_static_scratch (no allocation; feature tiles in a file-scope device array) against
ngpb_module_inference, through the C ABI. Collected last on purpose (zz).

The entry point has no counterpart in the reference.
Tolerance: the one test_mlp_gpu.py states for network outputs (2^-8 * max|.| + 2^-12). Both entry points launch
the same kernels on the same tiles, so the difference measured on B200 is zero; the bar is kept at the stated one.
"""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

IMAGE_ENC = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.26}
VOLUME_ENC = {"otype": "HashGrid", "n_levels": 8, "n_features_per_level": 4, "log2_hashmap_size": 15, "base_resolution": 16, "per_level_scale": 1.5}
NET = {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2}


def _static(ngp, module, x, out):
    """One call of the new entry point on torch's current stream; `out` is the caller's buffer."""
    lib = ngp._lib.lib
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    return lib.ngpb_module_inference_static_scratch(
        module._h, stream, x.shape[0], ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(module.params.data_ptr()))


def _assert_close(got, want):
    got = got.float().cpu().numpy()
    want = want.float().cpu().numpy()
    tol = 2.0 ** -8 * np.abs(want).max() + 2.0 ** -12
    err = np.abs(got - want).max()
    assert np.abs(want).max() > 1e-3, "degenerate reference output"
    assert err <= tol, (err, tol)
    return err


@pytest.mark.parametrize("dims,enc", [(2, IMAGE_ENC), (3, VOLUME_ENC)])
def test_static_scratch_inference_matches_allocating_entry_point(ngp, dims, enc):
    m = ngp.NetworkWithInputEncoding(dims, 3, enc, NET)
    g = torch.Generator(device="cpu").manual_seed(11 + dims)
    params = torch.empty(m.n_params, dtype=torch.float32).uniform_(-0.5, 0.5, generator=g)
    m.set_params(params.to(torch.float16).cuda())
    limit = int(ngp._lib.lib.ngpb_static_scratch_bytes())
    assert limit == 4 << 20
    n_max = limit // (32 * 2)                       # 32 encoded features, fp16
    # Two calls in a row on one stream: the largest batch the scratch holds, then a small one that reuses the
    # array while it still holds the first call's tiles.
    for n in (n_max, 1024):
        x = torch.rand((n, dims), generator=g, dtype=torch.float32).cuda()
        want = m.inference(x)
        out = torch.full((n, m.n_output_dims), float("nan"), dtype=torch.float16, device="cuda")
        rc = _static(ngp, m, x, out)
        assert rc == 0, ngp._lib.lib.ngpb_last_error().decode()
        torch.cuda.synchronize()
        assert torch.isfinite(out.float()).all()
        _assert_close(out, want)


def test_static_scratch_bare_network(ngp):
    """ngpb_create_network's module is a network behind an identity encoding: its input tiles are the scratch."""
    m = ngp.Network(32, 3, NET)
    g = torch.Generator(device="cpu").manual_seed(5)
    m.set_params(torch.empty(m.n_params, dtype=torch.float32).uniform_(-0.3, 0.3, generator=g).to(torch.float16).cuda())
    for n in (4096, 256):
        x = (torch.rand((n, 32), generator=g, dtype=torch.float32) * 2 - 1).cuda()
        want = m.inference(x)
        out = torch.full((n, m.n_output_dims), float("nan"), dtype=torch.float16, device="cuda")
        assert _static(ngp, m, x, out) == 0, ngp._lib.lib.ngpb_last_error().decode()
        torch.cuda.synchronize()
        _assert_close(out, want)


def test_static_scratch_rejects_what_it_cannot_hold(ngp):
    lib = ngp._lib.lib
    m = ngp.NetworkWithInputEncoding(2, 3, IMAGE_ENC, NET)
    m.set_params(torch.zeros(m.n_params, dtype=torch.float16, device="cuda"))
    n = (int(lib.ngpb_static_scratch_bytes()) // 64) + 256
    x = torch.zeros((n, 2), dtype=torch.float32, device="cuda")
    out = torch.zeros((n, m.n_output_dims), dtype=torch.float16, device="cuda")
    assert _static(ngp, m, x, out) != 0
    msg = lib.ngpb_last_error().decode()
    assert "static scratch holds" in msg and str(n) in msg
    # an encoding alone is not a network: not this entry point's module
    enc = ngp.Encoding(2, IMAGE_ENC)
    enc.set_params(torch.zeros(enc.n_params, dtype=torch.float16, device="cuda"))
    xe = torch.zeros((256, 2), dtype=torch.float32, device="cuda")
    oe = torch.zeros((256, enc.n_output_dims), dtype=torch.float16, device="cuda")
    assert _static(ngp, enc, xe, oe) != 0
    assert "NetworkWithInputEncoding" in lib.ngpb_last_error().decode()
    # a batch that is not a multiple of the granularity fails as it does in ngpb_module_inference
    x3 = torch.zeros((100, 2), dtype=torch.float32, device="cuda")
    o3 = torch.zeros((100, m.n_output_dims), dtype=torch.float16, device="cuda")
    assert _static(ngp, m, x3, o3) != 0
