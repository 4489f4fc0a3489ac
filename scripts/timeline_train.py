"""Kernel timeline of a few NeRF training steps (CUPTI activity trace through torch.profiler): which
kernels of which stream overlap. Diagnosis only - nothing measured here is a bench number.
Usage: python scripts/timeline_train.py [pretrain_steps] [traced_steps] [n_images] [width] [height]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import ProfilerActivity, profile

ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")

pre = int(sys.argv[1]) if len(sys.argv) > 1 else 300
traced = int(sys.argv[2]) if len(sys.argv) > 2 else 4
n_images = int(sys.argv[3]) if len(sys.argv) > 3 else 20
w = int(sys.argv[4]) if len(sys.argv) > 4 else 960
h = int(sys.argv[5]) if len(sys.argv) > 5 else 540
batch = 1 << 18

tb = ngp.Testbed()
scene, poses, focal = syn.make_dataset(tb, n_images=n_images, width=w, height=h, aabb_scale=4, amp_scale=8.0)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
for i in range(pre):
    tb.train(batch)
while tb.training_step % 16 != 1:   # start right after an occupancy-grid refresh
    tb.train(batch)
torch.cuda.synchronize()

with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(traced):
        tb.train(batch)
    torch.cuda.synchronize()

out = os.path.join(ROOT, "gpurun_out", "timeline_train.json")
os.makedirs(os.path.dirname(out), exist_ok=True)
prof.export_chrome_trace(out)

import json
ev = json.load(open(out))["traceEvents"]
ks = [e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "ts" in e]
ks.sort(key=lambda e: e["ts"])
t0 = ks[0]["ts"] if ks else 0
print(f"{len(ks)} device activities")
for e in ks:
    a = e.get("args", {})
    print(f"{e['ts'] - t0:10.1f} us  +{e['dur']:8.1f} us  stream {a.get('stream', '?'):>3}  grid {str(a.get('grid', '')):>16} {e['name'][:70]}")
