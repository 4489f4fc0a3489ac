"""Does a long-running, low-footprint co-resident kernel slow the training kernels, and which?
Phase timer of the training step (prefetch off) with and without a background kernel on another stream."""
import ctypes
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
tbm = importlib.import_module("instant-ngp-rendering_b200.testbed")
lib = ngp.lib

tb = ngp.Testbed()
tb.use_current_stream()
tb._set("train_prefetch", 0)
syn.make_dataset(tb, n_images=20, width=960, height=540, aabb_scale=4, amp_scale=8.0)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
for _ in range(300):
    tb.train(1 << 18)
table = torch.zeros(1 << 23, dtype=torch.int32, device="cuda")
sink = torch.zeros(4, dtype=torch.int32, device="cuda")
side = torch.cuda.Stream()


def phases(n=24):
    while tb.training_step % 16 != 1:
        tb.train(1 << 18)
    tbm.profile_enable(True)
    for _ in range(n):
        if tb.training_step % 16 == 0:
            break
        tb.train(1 << 18)
    torch.cuda.synchronize()
    p = tbm.profile_collect()
    tbm.profile_enable(False)
    return {k: ms / max(c, 1) for k, (ms, c) in p.items()}


base = phases()
cases = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]] or [(76, 128)]
for blocks, threads in cases:
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(side):
        e0.record(side)
        ngp._lib.check(lib.ngpb_probe_resident(ctypes.c_void_p(side.cuda_stream), blocks, threads, 60_000_000))   # 60 ms of residency
        e1.record(side)
    m0.record()
    with_bg = phases(12)
    m1.record()
    torch.cuda.synchronize()
    print(f"resident background {blocks} blocks x {threads} threads (sleeping): background ran {e0.elapsed_time(e1):.1f} ms, "
          f"measured steps started {e0.elapsed_time(m0):.2f} ms after it and took {m0.elapsed_time(m1):.1f} ms")
    for k in sorted(base):
        if k in with_bg:
            print(f"   {k:28s} {base[k] * 1e3:8.1f} us -> {with_bg[k] * 1e3:8.1f} us")
