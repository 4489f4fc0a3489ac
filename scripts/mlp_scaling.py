"""Kernel time against batch size for the NerfNetwork kernels (phase timer): what part of the training step's kernels is fixed cost.
Usage: python scripts/mlp_scaling.py   -> gpurun_out/mlp_scaling.json"""
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
tbm = importlib.import_module("instant-ngp-rendering_b200.testbed")
g = np.random.default_rng(1337)
m = ngp.NerfNetwork(rc.GRID_L8F4, rc.NERF_DIR, rc.NERF_DENSITY, rc.NERF_RGB)
params = np.concatenate([rc.mlp_weights(g, 32, 1), rc.mlp_weights(g, 32, 2), (g.standard_normal(rc.grid_n_params(3, rc.GRID_L8F4)) * 0.1).astype(np.float16)])
m.set_params(torch.from_numpy(params).cuda())
out = {}
for log2 in (14, 16, 17, 18, 19, 20, 22):
    n = 1 << log2
    x = g.random((n, 7), dtype=np.float32)
    o = g.random((n // 64, 1, 3), dtype=np.float32) * 0.5 + 0.1
    d = g.standard_normal((n // 64, 1, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=2, keepdims=True)
    x[:, :3] = np.clip(o + d * (np.arange(64, dtype=np.float32)[None, :, None] / 1024.0), 0, 0.999).reshape(-1, 3)
    dy = np.zeros((n, 16), dtype=np.float16)
    dy[:, :4] = (g.standard_normal((n, 4)) * 0.1).astype(np.float16)
    xt, dyt = torch.from_numpy(x).cuda(), torch.from_numpy(dy).cuda()
    y, ctx = m.forward(xt)
    for _ in range(3):
        m.inference(xt); m.backward(ctx, xt, y, dyt)
    torch.cuda.synchronize()
    tbm.profile_enable(True)
    reps = 10
    for _ in range(reps):
        m.inference(xt); m.backward(ctx, xt, y, dyt)
    torch.cuda.synchronize()
    prof = tbm.profile_collect()
    tbm.profile_enable(False)
    out[n] = {k: ms / cnt * 1e3 for k, (ms, cnt) in prof.items()}
    print(n, {k: round(v, 1) for k, v in out[n].items()})
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "mlp_scaling.json"), "w"), indent=1)
