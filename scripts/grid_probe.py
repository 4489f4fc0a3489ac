"""Measurements behind the hash-grid backward design (config 1: L16 F2 T=2^19, 2^20 uniform points):
  * random reduction rate with 4- / 8- / 16-byte vector reds, over tables of several sizes
  * L2 -> SM streaming bandwidth (the denominator of the sector-based roofline)
  * cost of the coarse levels: backward with all 16 levels against the same grid without its 2 / 3 coarsest
Usage: python scripts/grid_probe.py    -> gpurun_out/grid_probe.json"""
import ctypes
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

ngp = importlib.import_module("instant-ngp-rendering_b200")
lib = ngp._lib.lib
sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, reps=5, flush_l2=False):
    fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(reps):
        if flush_l2:
            flush.zero_()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record(); fn(); t1.record()
        torch.cuda.synchronize()
        ms.append(t0.elapsed_time(t1))
    return sorted(ms)[len(ms) // 2]


out = {}
table = torch.zeros(64 << 20, dtype=torch.uint8, device="cuda")
sink = torch.zeros(4, dtype=torch.int32, device="cuda")
tp = ctypes.c_void_p(table.data_ptr())
for log2 in (12, 19, 23):
    for vec in (1, 2, 4):
        ms = timed(lambda: lib.ngpb_probe_scatter_vec(sp, tp, log2, vec, 1 << 20, 128))
        out["red_v%d_2^%d" % (vec, log2)] = {"ms": ms, "g_instr_per_s": (1 << 20) * 128 / ms / 1e6, "gbytes_per_s": (1 << 20) * 128 * 4 * vec / ms / 1e6}
for mb in (16, 32, 64):
    passes = 8
    ms = timed(lambda: lib.ngpb_probe_l2_stream(sp, tp, mb << 20, passes, 148 * 8, ctypes.c_void_p(sink.data_ptr())))
    out["l2_stream_%dMB" % mb] = {"ms": ms, "gbytes_per_s": (mb << 20) * passes / ms / 1e6}

base = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.3819}
n = 1 << 20
g = torch.Generator(device="cuda").manual_seed(3)
x = torch.rand((n, 3), generator=g, device="cuda", dtype=torch.float32)
for name, cfg in (("L16", base), ("L14_from_level2", dict(base, n_levels=14, base_resolution=int(16 * 1.3819 ** 2 + 0.5))), ("L13_from_level3", dict(base, n_levels=13, base_resolution=int(16 * 1.3819 ** 3 + 0.5)))):
    enc = ngp.Encoding(3, cfg)
    enc.initialize_params()
    y, ctx = enc.forward(x)
    dy = (torch.randn(y.shape, generator=g, device="cuda", dtype=torch.float32) * 0.1).to(torch.float16)
    f_ms = timed(lambda: enc.inference(x), flush_l2=True)
    b_ms = timed(lambda: enc.backward(ctx, x, y, dy), flush_l2=True)
    out["grid_" + name] = {"levels": cfg["n_levels"], "base_resolution": cfg["base_resolution"], "fwd_ms": f_ms, "bwd_ms": b_ms, "bwd_us_per_level": b_ms * 1e3 / cfg["n_levels"]}
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "grid_probe.json"), "w"), indent=1)
for k, v in out.items():
    print(k, {a: (round(b, 4) if isinstance(b, float) else b) for a, b in v.items()})
