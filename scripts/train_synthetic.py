"""Trains the synthetic blob scene for a few hundred steps and prints loss / sample counters.
Usage: python scripts/train_synthetic.py [steps] [batch_log2] [aabb_scale]"""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
batch = 1 << (int(sys.argv[2]) if len(sys.argv) > 2 else 16)
aabb = int(sys.argv[3]) if len(sys.argv) > 3 else 1
tb = ngp.Testbed()
syn.make_dataset(tb, n_images=16, width=200, height=150, aabb_scale=aabb)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
t0 = time.time()
for i in range(steps):
    tb.train(batch)
    if i % 16 == 0 or i == steps - 1:
        print(f"step {tb.training_step:5d} loss {tb.loss:.5f} rays {tb.rays_per_batch:7d} samples {tb.measured_batch_size_before_compaction:8d} -> {tb.measured_batch_size:7d}", flush=True)
torch.cuda.synchronize()
print(f"{steps} steps in {time.time() - t0:.2f}s")
