"""Iterations of the training ray-march loop per ray (diagnostic build with -DNGPB_MARCH_STATS only:
  nvcc ... -DNGPB_MARCH_STATS -fmad=false -c nerf_train.cu, relink, run this, rebuild normally).
Usage: python scripts/march_stats.py [pretrain_steps] [n_images] [width] [height]"""
import ctypes
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
libm = importlib.import_module("instant-ngp-rendering_b200._lib")
lib = ctypes.CDLL(libm.LIB_PATH) if hasattr(libm, "LIB_PATH") else libm.lib

pre = int(sys.argv[1]) if len(sys.argv) > 1 else 300
n_images = int(sys.argv[2]) if len(sys.argv) > 2 else 20
w = int(sys.argv[3]) if len(sys.argv) > 3 else 960
h = int(sys.argv[4]) if len(sys.argv) > 4 else 540
tb = ngp.Testbed()
tb._set("train_prefetch", 0)
syn.make_dataset(tb, n_images=n_images, width=w, height=h, aabb_scale=4, amp_scale=8.0)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
for _ in range(pre):
    tb.train(1 << 18)
torch.cuda.synchronize()
hist = (ctypes.c_uint32 * 256)()
assert lib.ngpb_debug_march_hist(hist, 1) == 0
steps = 20
for _ in range(steps):
    tb.train(1 << 18)
torch.cuda.synchronize()
assert lib.ngpb_debug_march_hist(hist, 0) == 0
hcount = np.array(hist[:], dtype=np.int64)
n = hcount.sum()
centers = np.arange(256) * 16 + 8
print(f"rays/step {n / steps:.0f}  rays_per_batch {tb.rays_per_batch}  samples/step {tb.measured_batch_size_before_compaction}")
print(f"mean iterations {float((hcount * centers).sum()) / n:.1f}")
cum = np.cumsum(hcount) / n
for q in (0.5, 0.9, 0.99, 0.999, 1.0):
    print(f"  p{q * 100:g}: {centers[np.searchsorted(cum, q - 1e-12)]} iterations")
print("histogram (16 per bin):", hcount[: np.max(np.nonzero(hcount)) + 1].tolist())
