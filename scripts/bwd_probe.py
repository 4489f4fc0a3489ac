"""Times NetworkWithInputEncoding.backward / Encoding.backward on unordered points (image shape 2-D L16 F2, config 1 3-D L16 F2).
Usage: [NGPB_GRID_COARSE=0] python scripts/bwd_probe.py"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

ngp = importlib.import_module("instant-ngp-rendering_b200")
cfgs = importlib.import_module("baseline.configs")


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(reps):
        fn()
    t1.record()
    torch.cuda.synchronize()
    return t0.elapsed_time(t1) / reps


g = torch.Generator(device="cuda").manual_seed(3)
n = 1 << 18
img = cfgs.IMAGE_BASE_CONFIG
for name, dims, enc in (("image_2d", 2, img["encoding"]), ("config1_3d", 3, {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19, "base_resolution": 16, "per_level_scale": 1.3819})):
    e = ngp.Encoding(dims, enc)
    e.initialize_params()
    x = torch.rand((n, dims), generator=g, device="cuda", dtype=torch.float32)
    y, ctx = e.forward(x)
    dy = (torch.randn(y.shape, generator=g, device="cuda", dtype=torch.float32) * 0.1).to(torch.float16)
    print(name, "levels", e.level_table()[2], "fwd %.4f ms" % timed(lambda: e.inference(x)), "bwd %.4f ms" % timed(lambda: e.backward(ctx, x, y, dy)), "coarse", os.environ.get("NGPB_GRID_COARSE", "1"))
m = ngp.NetworkWithInputEncoding(2, 3, img["encoding"], img["network"])
m.initialize_params()
x = torch.rand((n, 2), generator=g, device="cuda", dtype=torch.float32)
y, ctx = m.forward(x)
dy = (torch.randn(y.shape, generator=g, device="cuda", dtype=torch.float32) * 0.1).to(torch.float16)
print("image network: fwd %.4f ms, bwd %.4f ms" % (timed(lambda: m.inference(x)), timed(lambda: m.backward(ctx, x, y, dy))))
