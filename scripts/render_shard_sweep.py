"""Window-policy sweep for ONE of 8 tile shards of a 1920x1080 frame of the bench scene (what a rank of the
8-GPU render does), on one GPU. Usage: python scripts/render_shard_sweep.py [train_steps]"""
import importlib
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 600
tb = ngp.Testbed()
scene, poses, focal = syn.make_dataset(tb, n_images=50, width=1920, height=1080, aabb_scale=4, amp_scale=8.0)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
for i in range(steps):
    tb.train(1 << 18)
view = syn.look_at_ngp(np.array([0.5, 0.5, 0.5]) + 1.4 * np.array([0.62, -0.55, 0.56]))
policies = [(8 << 20, 4, 32, 4), (8 << 20, 4, 32, 8), (8 << 20, 4, 16, 4), (8 << 20, 4, 24, 4), (6 << 20, 4, 24, 4), (4 << 20, 4, 16, 4), (4 << 20, 4, 16, 8),
            (12 << 20, 4, 32, 4), (8 << 20, 2, 32, 4), (8 << 20, 8, 32, 4), (3 << 20, 4, 12, 8), (2 << 20, 4, 8, 8)]
for shard in ((0, 8), (3, 8), (0, 1)):
    for policy in policies:
        tb._set("render_target_queries", policy[0]); tb._set("render_min_steps", policy[1]); tb._set("render_max_steps", policy[2]); tb._set("render_windows_per_sync", policy[3])
        for _ in range(2):
            tb.render_nerf(1920, 1080, view, focal, shard=shard)
        reps = 10
        torch.cuda.synchronize()
        t0 = time.time()
        for _ in range(reps):
            rgba, depth = tb.render_nerf(1920, 1080, view, focal, shard=shard)
        torch.cuda.synchronize()
        dt = (time.time() - t0) / reps
        print(f"shard {shard} policy ({policy[0] >> 20}M, {policy[1]}, {policy[2]}, {policy[3]}): {dt * 1e3:.3f} ms/frame, {tb.last_render_samples / (1920 * 1080 / shard[1]):.1f} queries/ray", flush=True)
