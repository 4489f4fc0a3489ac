"""Wall time per NeRF training step (CUDA events around K Testbed.train calls), prefetch on and off.
Usage: [NGPB_MARCH_WARPS=n] python scripts/step_time.py [pretrain] [steps] [n_images] [width] [height]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")

pre = int(sys.argv[1]) if len(sys.argv) > 1 else 300
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
n_images = int(sys.argv[3]) if len(sys.argv) > 3 else 20
w = int(sys.argv[4]) if len(sys.argv) > 4 else 960
h = int(sys.argv[5]) if len(sys.argv) > 5 else 540
tb = ngp.Testbed()
tb.use_current_stream()
syn.make_dataset(tb, n_images=n_images, width=w, height=h, aabb_scale=4, amp_scale=8.0)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
for _ in range(pre):
    tb.train(1 << 18)
for prefetch in (1, 0, 1, 0):
    tb._set("train_prefetch", prefetch)
    for _ in range(20):
        tb.train(1 << 18)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 0
    t0.record()
    for _ in range(steps):
        tb.train(1 << 18)
        n += tb.measured_batch_size
    t1.record()
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / steps
    print(f"warps {os.environ.get('NGPB_MARCH_WARPS', 'default')} prefetch {prefetch}: {ms:.4f} ms/step, {n / steps / ms / 1e3:.1f} M samples/s, rays {tb.rays_per_batch}", flush=True)
