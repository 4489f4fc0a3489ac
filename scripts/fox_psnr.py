"""The reference's own fox dataset (data/nerf/fox: 50 JPEGs 1080x1920, OpenCV lens, aabb_scale 4) through the pyngp
shell: load_training_data -> configs/nerf/base.json -> train -> scripts/run.py's PSNR protocol on the training
views (the dataset ships no test split). The data is staged under data_stage/ (git-ignored; copied from
/root/reference/data in the authoring container) so that it travels to the GPU box; nothing is committed.

Usage: python scripts/fox_psnr.py [data_stage/fox] [steps,steps,...]    -> gpurun_out/fox_psnr.json
"""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "instant-ngp-rendering_b200"))
import numpy as np
import torch
import pyngp as ngp

syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")

scene = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "data_stage", "fox")
checkpoints = [int(s) for s in (sys.argv[2] if len(sys.argv) > 2 else "1000,2000,5000").split(",")]

testbed = ngp.Testbed()
t0 = time.time()
testbed.load_training_data(scene)
load_s = time.time() - t0
testbed.reload_network_from_json(dict(syn.NERF_BASE_CONFIG))     # = the reference's configs/nerf/base.json (this fork: L8 F4)
testbed.nerf.training.near_distance = 0.1                        # scripts/run.py default for real scenes
testbed.shall_train = True
n_images = testbed.nerf.training.dataset.n_images


def evaluate(stride=5, spp=4):
    """scripts/run.py:216-268 on every `stride`-th training view."""
    testbed.background_color = [0.0, 0.0, 0.0, 1.0]
    testbed.snap_to_pixel_centers = True
    testbed.nerf.render_min_transmittance = 1e-4
    testbed.shall_train = False
    psnrs = []
    for i in range(0, n_images, stride):
        resolution = testbed.nerf.training.dataset.metadata[i].resolution
        testbed.render_ground_truth = True
        testbed.set_camera_to_training_view(i)
        ref_image = testbed.render(resolution[0], resolution[1], 1, True)
        testbed.render_ground_truth = False
        image = testbed.render(resolution[0], resolution[1], spp, True)
        # run.py compares in sRGB after clipping (common.py compute_error "MSE" on linear_to_srgb images)
        to_srgb = lambda x: np.where(x < 0.0031308, 12.92 * x, 1.055 * np.maximum(x, 1e-9) ** (1 / 2.4) - 0.055)
        a, b = np.clip(to_srgb(image[..., :3]), 0, 1), np.clip(to_srgb(ref_image[..., :3]), 0, 1)
        mse = float(np.mean((a - b) ** 2))
        psnrs.append(-10.0 * np.log10(max(mse, 1e-12)))
    testbed.shall_train = True
    testbed.snap_to_pixel_centers = False
    testbed.nerf.render_min_transmittance = 0.01
    return float(np.mean(psnrs)), float(np.min(psnrs)), float(np.max(psnrs))


out = {"scene": scene, "n_images": n_images, "load_s": load_s, "aabb_scale": testbed.nerf.training.dataset.aabb_scale, "checkpoints": []}
done = 0
for target in checkpoints:
    torch.cuda.synchronize()
    t0 = time.time()
    samples = 0
    while testbed.training_step < target:
        testbed.frame()
        samples += testbed.measured_batch_size
    _ = testbed.loss
    torch.cuda.synchronize()
    dt = time.time() - t0
    steps = target - done
    done = target
    mean, lo, hi = evaluate()
    rec = {"step": target, "loss": testbed.loss, "train_s": dt, "ms_per_step": 1e3 * dt / steps, "samples_per_s": samples / dt,
           "psnr_mean": mean, "psnr_min": lo, "psnr_max": hi, "rays_per_batch": testbed.rays_per_batch}
    print(rec, flush=True)
    out["checkpoints"].append(rec)
# configs[3]: the trained fox at 1920x1080 from training view 0's pose (landscape frame, fov from the view) and from a
# pose between two views, 1 spp, render_min_transmittance 0.01: Mrays/s by CUDA events around render_nerf
testbed.set_camera_to_training_view(0)
cam_a = np.array(testbed._camera, dtype=np.float32)
testbed.set_camera_to_training_view(1)
cam_b = np.array(testbed._camera, dtype=np.float32)
mid = cam_a.copy()
mid[:, 3] = 0.5 * (cam_a[:, 3] + cam_b[:, 3])
rw, rh = 1920, 1080
focal = float(np.broadcast_to(testbed._relative_focal_length, (2,))[1]) * rh
out["render"] = []
for name, cam in (("training view 0", cam_a), ("between views 0 and 1", mid)):
    for _ in range(3):
        testbed.render_nerf(rw, rh, cam, focal)
    torch.cuda.synchronize()
    frames = 10
    t0 = time.time()
    for _ in range(frames):
        rgba, depth = testbed.render_nerf(rw, rh, cam, focal)
    torch.cuda.synchronize()
    ms = 1e3 * (time.time() - t0) / frames
    rec = {"view": name, "resolution": [rw, rh], "ms_per_frame": ms, "mrays_per_s": rw * rh / ms / 1e3, "queries_per_ray": testbed.last_render_samples / (rw * rh),
           "coverage": float((rgba[..., 3] > 0.5).float().mean())}
    print(rec, flush=True)
    out["render"].append(rec)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "fox_psnr.json"), "w") as f:
    json.dump(out, f, indent=1)
try:
    testbed.save_snapshot(os.path.join(ROOT, "gpurun_out", "fox_step%d.ingp" % done), False)
    print("snapshot bytes", os.path.getsize(os.path.join(ROOT, "gpurun_out", "fox_step%d.ingp" % done)))
except Exception as e:   # the snapshot is a by-product here
    print("snapshot failed:", e)
