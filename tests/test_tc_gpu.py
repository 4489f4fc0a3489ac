"""GPU tests of the tcgen05 building blocks: the three shared-memory operand views used by the
fused MLP kernels, against torch fp32 matmul of the same fp16 inputs (tolerance 1e-3 relative:
fp32 accumulation, products exact)."""
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _run(ngp, mode, K, N, X, W, rows_out):
    D = torch.zeros((rows_out, N), dtype=torch.float32, device="cuda")
    ngp._lib.check(ngp.lib.ngpb_selftest_umma(None, mode, K, N, X.data_ptr(), W.data_ptr(), D.data_ptr()))
    torch.cuda.synchronize()
    return D


@pytest.mark.parametrize("K,N", [(32, 64), (64, 64), (64, 16), (16, 32)])
def test_x_wT(ngp, K, N):
    g = torch.Generator(device="cuda").manual_seed(K * 100 + N)
    X = torch.randn((128, K), generator=g, device="cuda").half()
    W = torch.randn((N, K), generator=g, device="cuda").half()
    D = _run(ngp, 0, K, N, X, W, 128)
    ref = X.float() @ W.float().t()
    assert torch.allclose(D, ref, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("K,N", [(64, 32), (64, 64), (16, 64), (16, 32)])
def test_dy_w(ngp, K, N):
    g = torch.Generator(device="cuda").manual_seed(K * 100 + N + 1)
    X = torch.randn((128, K), generator=g, device="cuda").half()
    W = torch.randn((K, N), generator=g, device="cuda").half()
    D = _run(ngp, 1, K, N, X, W, 128)
    ref = X.float() @ W.float()
    assert torch.allclose(D, ref, rtol=1e-3, atol=1e-3)


@pytest.mark.parametrize("N", [16, 32, 64])
def test_pT_q(ngp, N):
    g = torch.Generator(device="cuda").manual_seed(N + 7)
    P = torch.randn((128, 64), generator=g, device="cuda").half()
    Q = torch.randn((128, N), generator=g, device="cuda").half()
    D = _run(ngp, 2, 64, N, P, Q, 64)
    ref = P.float().t() @ Q.float()
    assert torch.allclose(D, ref, rtol=1e-3, atol=2e-3)
