"""One protocol, two implementations: drives either the reference's OWN Python binding (baseline/_ref/pyngp*.so,
the unmodified src/python_api.cu built by baseline/build_ref.sh) or this repository's mirror of it
(instant-ngp-rendering_b200/pyngp.py over libngp_b200.so) through the same calls, the way scripts/run.py of the
reference drives a Testbed. One process per implementation (both modules are called `pyngp`).

    python baseline/run_pyngp.py --impl reference|ours --what fox[,albert][,synthetic] [--steps 1000,2000,5000]
        [--save PATH.ingp] [--load PATH.ingp] --out gpurun_out/NAME.json

fox        data_stage/fox (the reference's data/nerf/fox), configs/nerf/base.json, 2^18 samples per step:
           ms/step between checkpoints by wall clock (device synchronised), then scripts/run.py:216-268's PSNR protocol
           on every fifth training view, then the 1920x1080 1-spp render of training view 0's pose through render().
albert     data_stage/albert.exr (data/image/albert.exr), configs/image/base.json, 1000 + 1000 steps of 2^18 samples.
synthetic  bench.py's synthetic fox-shaped dataset through create_empty_nerf_dataset / set_image (float images).
--load     skips training: load_snapshot(PATH), then the PSNR protocol (cross-implementation snapshot check).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "baseline"))
from configs import NERF_BASE_CONFIG, IMAGE_BASE_CONFIG  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--impl", choices=["reference", "ours"], required=True)
ap.add_argument("--what", default="fox")
ap.add_argument("--steps", default="1000,2000,5000")
ap.add_argument("--save", default="")
ap.add_argument("--load", default="")
ap.add_argument("--out", default="")
ap.add_argument("--batch", type=int, default=1 << 18)
ap.add_argument("--eval", type=int, default=1, help="0: skip the PSNR protocol (timing only)")
ap.add_argument("--frames", type=int, default=10)
ap.add_argument("--native", type=int, default=0, help="ours only: 1 = scene, snapshots through the library's host C++ (ngpb_testbed_load_training_data / save_snapshot / load_snapshot)")
args = ap.parse_args()

sys.path.insert(0, os.path.join(ROOT, "baseline", "_ref") if args.impl == "reference" else os.path.join(ROOT, "instant-ngp-rendering_b200"))
import pyngp as ngp  # noqa: E402
import torch  # noqa: E402  (device synchronisation and nothing else)


def sync():
    torch.cuda.synchronize()


def to_srgb(x):
    return np.where(x < 0.0031308, 12.92 * x, 1.055 * np.maximum(x, 1e-9) ** (1 / 2.4) - 0.055)


def evaluate(testbed, stride=5, spp=4):
    """scripts/run.py:216-268 on every `stride`-th training view (fox ships no test split)."""
    testbed.background_color = [0.0, 0.0, 0.0, 1.0]
    testbed.snap_to_pixel_centers = True
    testbed.nerf.render_min_transmittance = 1e-4
    testbed.shall_train = False
    psnrs = []
    n_images = testbed.nerf.training.dataset.n_images
    for i in range(0, n_images, stride):
        resolution = [int(v) for v in testbed.nerf.training.dataset.metadata[i].resolution]
        testbed.render_ground_truth = True
        testbed.set_camera_to_training_view(i)
        ref_image = np.array(testbed.render(resolution[0], resolution[1], 1, True))
        testbed.render_ground_truth = False
        image = np.array(testbed.render(resolution[0], resolution[1], spp, True))
        a, b = np.clip(to_srgb(image[..., :3]), 0, 1), np.clip(to_srgb(ref_image[..., :3]), 0, 1)
        psnrs.append(-10.0 * np.log10(max(float(np.mean((a - b) ** 2)), 1e-12)))
    testbed.shall_train = True
    testbed.snap_to_pixel_centers = False
    testbed.nerf.render_min_transmittance = 0.01
    return {"psnr_mean": float(np.mean(psnrs)), "psnr_min": float(np.min(psnrs)), "psnr_max": float(np.max(psnrs)), "views": len(psnrs)}


def time_render(testbed, frames=None):
    frames = frames or args.frames
    testbed.shall_train = False
    testbed.set_camera_to_training_view(0)
    for _ in range(3):
        testbed.render(1920, 1080, 1, True)
    sync()
    t0 = time.time()
    for _ in range(frames):
        img = testbed.render(1920, 1080, 1, True)
    sync()
    ms = 1e3 * (time.time() - t0) / frames
    testbed.shall_train = True
    return {"resolution": [1920, 1080], "ms_per_frame_e2e": ms, "mrays_per_s_e2e": 1920 * 1080 / ms / 1e3, "coverage": float((np.array(img)[..., 3] > 0.5).mean())}


def run_fox(out):
    scene = os.path.join(ROOT, "data_stage", "fox")
    testbed = ngp.Testbed()
    testbed.root_dir = ROOT
    t0 = time.time()
    native = bool(args.native) and args.impl == "ours"
    (testbed.load_training_data_native if native else testbed.load_training_data)(scene)
    rec = {"scene": scene, "host": "c++" if native else "python binding", "load_s": time.time() - t0, "n_images": int(testbed.nerf.training.dataset.n_images), "checkpoints": []}
    if args.load:
        (testbed.load_snapshot_native if native else testbed.load_snapshot)(args.load)
        rec["loaded"] = args.load
        rec["training_step"] = int(testbed.training_step)
        rec["eval"] = evaluate(testbed)
        rec["render"] = time_render(testbed)
        out["fox"] = rec
        print(rec, flush=True)
        return
    testbed.reload_network_from_json(dict(NERF_BASE_CONFIG))
    testbed.nerf.training.near_distance = 0.1   # scripts/run.py default for real scenes
    testbed.shall_train = True
    testbed.training_batch_size = args.batch
    done = 0
    for target in [int(s) for s in args.steps.split(",")]:
        sync()
        t0 = time.time()
        while testbed.training_step < target:
            testbed.frame()
        sync()
        dt = time.time() - t0
        c = {"step": target, "loss": float(testbed.loss), "train_s": dt, "ms_per_step": 1e3 * dt / (target - done)}
        if hasattr(testbed, "measured_batch_size"):
            c["measured_batch_size"] = int(testbed.measured_batch_size)
        done = target
        if args.eval:
            c.update(evaluate(testbed))
        print(c, flush=True)
        rec["checkpoints"].append(c)
    rec["render"] = time_render(testbed)
    print(rec["render"], flush=True)
    if args.save:
        (testbed.save_snapshot_native if native else testbed.save_snapshot)(args.save, False)
        rec["saved"] = args.save
        rec["snapshot_bytes"] = os.path.getsize(args.save)
    out["fox"] = rec
    # samples per second over the last segment: the step trains min(measured_batch_size, batch) samples (the compacted
    # batch is clamped to the batch size, src/testbed_nerf.cu:2862-2870); the reference's binding does not expose the
    # counter, so its runs are credited with the full batch, which is an upper bound for it
    last = rec["checkpoints"][-1]
    mb = min(int(last.get("measured_batch_size", args.batch)), args.batch)
    rec["samples_per_s"] = mb / (last["ms_per_step"] * 1e-3)


def run_albert(out):
    path = os.path.join(ROOT, "data_stage", "albert.exr")
    testbed = ngp.Testbed()
    testbed.root_dir = ROOT
    testbed.load_training_data(path)
    testbed.reload_network_from_json(dict(IMAGE_BASE_CONFIG))
    testbed.shall_train = True
    testbed.training_batch_size = args.batch
    rec = {"image": path, "checkpoints": []}
    for lo, hi in ((0, 1000), (1000, 2000)):
        sync()
        t0 = time.time()
        while testbed.training_step < hi:
            testbed.frame()
        sync()
        dt = time.time() - t0
        c = {"step": hi, "loss": float(testbed.loss), "ms_per_step": 1e3 * dt / (hi - lo), "samples_per_s": (hi - lo) * args.batch / dt}
        print(c, flush=True)
        rec["checkpoints"].append(c)
    out["albert"] = rec


def main():
    out = {"impl": args.impl, "batch": args.batch}
    for what in args.what.split(","):
        {"fox": run_fox, "albert": run_albert}[what](out)
    if args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        with open(args.out, "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
