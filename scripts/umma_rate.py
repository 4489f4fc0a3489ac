"""Cycles per tcgen05.mma for the three operand views the fused MLP kernels use (tc_device.cuh), on one SM."""
import ctypes, importlib, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
ngp = importlib.import_module("instant-ngp-rendering_b200")
out = torch.zeros(2, dtype=torch.int64, device="cuda")
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
for mode, name, shapes in ((0, "X W^T (K-major A, K-major B)", [(64, 64), (32, 64), (64, 16)]), (1, "dY W (B MN-major)", [(64, 64), (16, 64), (64, 32)]), (2, "P^T Q (both MN-major, M=64)", [(64, 64), (64, 32), (64, 16)])):
    for K, N in shapes:
        for reps in (1, 16, 64):
            ngp._lib.check(ngp.lib.ngpb_probe_umma_rate(st, mode, K, N, reps, ctypes.c_void_p(out.data_ptr())))
            torch.cuda.synchronize()
            tot, iss = out.cpu().tolist()
            n_mma = reps * ((K // 16) if mode < 2 else 8)
            print(f"mode {mode} {name:32s} K={K} N={N} reps={reps:3d}: {n_mma:4d} MMAs, total {tot:7d} cyc ({tot / n_mma:7.1f}/MMA), issue {iss:6d} cyc ({iss / n_mma:6.1f}/MMA)")
