"""configs[1] on the reference's own image (data/image/albert.exr, 1024 x 1024) through the pyngp shell:
load_training_data(image file) -> image mode, configs/image/base.json, 1000 warm-up + 1000 timed steps of
2^18 samples, then the MSE / PSNR of the fitted image at the pixel centres. Data staged under data_stage/.

Usage: python scripts/albert_mse.py [data_stage/albert.exr]    -> gpurun_out/albert_mse.json
"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "instant-ngp-rendering_b200"))
import numpy as np
import torch
import pyngp as ngp
from bench import IMAGE_CONFIG

path = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "data_stage", "albert.exr")
testbed = ngp.Testbed()
testbed.load_training_data(path)
assert testbed.mode == ngp.TestbedMode.Image
testbed.reload_network_from_json(dict(IMAGE_CONFIG))
loader = __import__("importlib").import_module("instant-ngp-rendering_b200.loader")
target = torch.from_numpy(loader.load_image_file(path)).cuda()
h, w = target.shape[:2]
out = {"image": path, "resolution": [w, h], "checkpoints": []}
for lo, hi in ((0, 1000), (1000, 2000)):
    torch.cuda.synchronize()
    t0 = time.time()
    for _ in range(lo, hi):
        testbed.train(1 << 18)
    _ = testbed.loss
    torch.cuda.synchronize()
    dt = time.time() - t0
    got = testbed.render_image(w, h)
    mse = float(((got[..., :3] - target[..., :3]) ** 2).mean())
    rec = {"step": hi, "loss": testbed.loss, "ms_per_step": 1e3 * dt / (hi - lo), "samples_per_s": (hi - lo) * (1 << 18) / dt, "mse_linear": mse, "psnr_linear": -10 * np.log10(mse)}
    print(rec, flush=True)
    out["checkpoints"].append(rec)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
with open(os.path.join(ROOT, "gpurun_out", "albert_mse.json"), "w") as f:
    json.dump(out, f, indent=1)
