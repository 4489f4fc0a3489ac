"""Phase breakdown of the fox-shaped NeRF training step and of a 1080p render (library phase timer).
Usage: python scripts/profile_train.py [pretrain_steps] [timed_steps] [n_images] [width] [height]"""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
tbm = importlib.import_module("instant-ngp-rendering_b200.testbed")

pre = int(sys.argv[1]) if len(sys.argv) > 1 else 300
timed = int(sys.argv[2]) if len(sys.argv) > 2 else 50
n_images = int(sys.argv[3]) if len(sys.argv) > 3 else 50
w = int(sys.argv[4]) if len(sys.argv) > 4 else 1920
h = int(sys.argv[5]) if len(sys.argv) > 5 else 1080
batch = 1 << 18

tb = ngp.Testbed()
t0 = time.time()
scene, poses, focal = syn.make_dataset(tb, n_images=n_images, width=w, height=h, aabb_scale=4, amp_scale=8.0)
torch.cuda.synchronize()
print(f"dataset {n_images}x{w}x{h} in {time.time() - t0:.1f}s", flush=True)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
t0 = time.time()
for i in range(pre):
    tb.train(batch)
    if i % 50 == 0:
        print(f"step {tb.training_step:5d} loss {tb.loss:.5f} rays {tb.rays_per_batch:7d} samples {tb.measured_batch_size_before_compaction:8d} -> {tb.measured_batch_size:7d}", flush=True)
print(f"pretrain {pre} steps in {time.time() - t0:.2f}s", flush=True)

tbm.profile_enable(True)
t0 = time.time()
n_samples = 0
n_before = 0
for i in range(timed):
    tb.train(batch)
    n_samples += tb.measured_batch_size
    n_before += tb.measured_batch_size_before_compaction
dt = time.time() - t0
prof = tbm.profile_collect()
tbm.profile_enable(False)
print(f"{timed} steps in {dt * 1e3:.1f} ms = {dt / timed * 1e3:.3f} ms/step, {n_samples / dt / 1e6:.1f} M compacted samples/s, {n_before / dt / 1e6:.1f} M marched samples/s; loss {tb.loss:.5f}")
for k, (ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print(f"  {k:32s} {ms / timed:8.4f} ms/step  ({cnt} intervals)")

for policy in ((2 << 20, 1, 8, 8), (8 << 20, 2, 8, 8), (16 << 20, 4, 8, 16), (16 << 20, 8, 8, 16)):
    tb._set("render_target_queries", policy[0]); tb._set("render_min_steps", policy[1]); tb._set("render_max_steps", policy[2]); tb._set("render_windows_per_sync", policy[3])
    tb.render_nerf(1920, 1080, poses[0], focal * 1920 / w)
    tbm.profile_enable(True)
    t0 = time.time()
    reps = 5
    for _ in range(reps):
        rgba, depth = tb.render_nerf(1920, 1080, poses[0], focal * 1920 / w)
    torch.cuda.synchronize()
    dt = (time.time() - t0) / reps
    prof = tbm.profile_collect()
    tbm.profile_enable(False)
    print(f"render policy {policy}: {dt * 1e3:.2f} ms/frame = {1920 * 1080 / dt / 1e6:.1f} Mrays/s, {tb.last_render_samples / (1920 * 1080):.1f} queries/ray, coverage {float((rgba[..., 3] > 0.5).float().mean()):.2f}")
    for k, (ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
        print(f"  {k:32s} {ms / reps:8.4f} ms/frame  ({cnt // reps} intervals/frame)")
