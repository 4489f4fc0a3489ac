"""Times NerfNetwork inference / forward+backward at the benchmark shape (2^18 samples, base.json L8F4 T=2^19) with CUDA
events and an L2 flush between iterations: the measurement baseline/harness.cu makes of the reference's NerfNetwork
(profiles/r02_reference_kernel_timings.json: 0.157 ms inference, 0.827 ms fwd+bwd). Env: NGPB_FUSED=0 selects the
unfused kernels, NGPB_FUSED_GROUPS=g caps the worker groups per CTA."""
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_gpu_cases as rc

ngp = importlib.import_module("instant-ngp-rendering_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
ray_coherent = len(sys.argv) > 2 and sys.argv[2] == "rays"
g = np.random.default_rng(1337)
m = ngp.NerfNetwork(rc.GRID_L8F4, rc.NERF_DIR, rc.NERF_DENSITY, rc.NERF_RGB)
params = np.concatenate([rc.mlp_weights(g, 32, 1), rc.mlp_weights(g, 32, 2), (g.standard_normal(rc.grid_n_params(3, rc.GRID_L8F4)) * 0.1).astype(np.float16)])
m.set_params(torch.from_numpy(params).cuda())
x = g.random((n, 7), dtype=np.float32)
if ray_coherent:   # 64 consecutive samples share a ray: positions advance 1/1024 per sample
    o = g.random((n // 64, 1, 3), dtype=np.float32) * 0.5 + 0.1
    d = g.standard_normal((n // 64, 1, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=2, keepdims=True)
    x[:, :3] = np.clip(o + d * (np.arange(64, dtype=np.float32)[None, :, None] / 1024.0), 0, 0.999).reshape(-1, 3)
dy = np.zeros((n, 16), dtype=np.float16)
dy[:, :4] = (g.standard_normal((n, 4)) * 0.1).astype(np.float16)
xt, dyt = torch.from_numpy(x).cuda(), torch.from_numpy(dy).cuda()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, iters=20):
    ms = []
    for i in range(iters + 3):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        if i >= 3:
            ms.append(a.elapsed_time(b))
    return float(np.median(ms))


out, ctx = m.forward(xt)
inf = timed(lambda: m.inference(xt))
bwd = timed(lambda: m.backward(ctx, xt, out, dyt))
print(json.dumps({"n": n, "points": "ray-coherent" if ray_coherent else "uniform", "fused": os.environ.get("NGPB_FUSED", "1"), "groups": os.environ.get("NGPB_FUSED_GROUPS", "auto"),
                  "inference_ms": inf, "fwd_bwd_scatter_ms": bwd}))
