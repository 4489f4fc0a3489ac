"""Phase breakdown of a 1920x1080 render of the bench scene, whole frame and one of 8 tile shards.
Usage: python scripts/profile_render.py [train_steps]"""
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
tbm = importlib.import_module("instant-ngp-rendering_b200.testbed")

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 600
fox = len(sys.argv) > 2 and sys.argv[2] == "fox"
tb = ngp.Testbed()
if fox:
    tb.load_training_data(os.path.join(ROOT, "data_stage", "fox"))
    tb.nerf.training.near_distance = 0.1
else:
    scene, poses, focal = syn.make_dataset(tb, n_images=50, width=1920, height=1080, aabb_scale=4, amp_scale=8.0)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
for i in range(steps):
    tb.train(1 << 18)
if fox:
    tb.set_camera_to_training_view(0)
    cam_a = np.array(tb._camera, dtype=np.float32)
    tb.set_camera_to_training_view(1)
    view = cam_a.copy()
    view[:, 3] = 0.5 * (cam_a[:, 3] + np.array(tb._camera, dtype=np.float32)[:, 3])
    focal = float(np.broadcast_to(tb._relative_focal_length, (2,))[1]) * 1080
else:
    view = syn.look_at_ngp(np.array([0.5, 0.5, 0.5]) + 1.4 * np.array([0.62, -0.55, 0.56]))
policies = {(0, 1): ((12 << 20, 4, 32, 4), (24 << 20, 4, 32, 4), (48 << 20, 8, 32, 4)),
            (0, 8): ((12 << 17, 4, 32, 4), (12 << 18, 4, 32, 4), (12 << 19, 4, 32, 4), (12 << 20, 4, 32, 4), (24 << 20, 4, 32, 4), (24 << 20, 8, 64, 4), (24 << 20, 16, 64, 8))}
for shard in ((0, 1), (0, 8)):
    for policy in policies[shard]:
        tb._set("render_target_queries", policy[0]); tb._set("render_min_steps", policy[1]); tb._set("render_max_steps", policy[2]); tb._set("render_windows_per_sync", policy[3])
        tb.render_nerf(1920, 1080, view, focal, shard=shard)
        tbm.profile_enable(True)
        reps = 5
        torch.cuda.synchronize()
        t0 = time.time()
        for _ in range(reps):
            rgba, depth = tb.render_nerf(1920, 1080, view, focal, shard=shard)
        torch.cuda.synchronize()
        dt = (time.time() - t0) / reps
        prof = tbm.profile_collect()
        tbm.profile_enable(False)
        print(f"shard {shard} policy {policy}: {dt * 1e3:.2f} ms/frame, {tb.last_render_samples / (1920 * 1080 / shard[1]):.1f} queries/ray")
        print("   " + ", ".join(f"{k} {ms / reps:.2f} ({cnt // reps})" for k, (ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0])))
