"""Loss trajectories of the large-table case (T = 2^22, aabb_scale 4): one testbed against two emulated
shards, with and without the prefetch. Diagnosis tool for tests/test_dp_gpu.py."""
import importlib
import sys

import numpy as np
import torch

ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
check, lib = ngp._lib.check, ngp.lib
cfg = dict(syn.NERF_BASE_CONFIG)
cfg["encoding"] = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 22, "base_resolution": 16}
batch = 1 << 16
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 145


def make(rank=0, world=1, prefetch=True):
    tb = ngp.Testbed()
    tb.use_current_stream()
    syn.make_dataset(tb, n_images=8, width=96, height=72, aabb_scale=4)
    tb.reload_network_from_json(cfg)
    tb._set("train_prefetch", int(prefetch))
    if world > 1:
        tb.set_shard(rank, world)
    return tb


def run_single(prefetch):
    tb = make(prefetch=prefetch)
    out = []
    for step in range(steps):
        tb.train(batch)
        if step % 16 == 0:
            out.append((tb.loss, tb.measured_batch_size, tb.rays_per_batch))
    tb.close()
    return out


def run_sharded():
    tbs = [make(r, 2) for r in range(2)]
    grads = [tb.device_buffer("gradients", torch.float16) for tb in tbs]
    counters = [tb.device_buffer("counters", torch.int32) for tb in tbs]
    out = []
    for step in range(steps):
        for tb in tbs:
            check(lib.ngpb_testbed_train_forward(tb._h, batch))
        cs = counters[0][:3] + counters[1][:3]
        ls = counters[0][3:4].view(torch.float32) + counters[1][3:4].view(torch.float32)
        for k in range(2):
            counters[k][:3].copy_(cs)
            counters[k][3:4].view(torch.float32).copy_(ls)
        for tb in tbs:
            check(lib.ngpb_testbed_train_backward(tb._h, batch))
        g = grads[0].float() + grads[1].float()
        for k in range(2):
            grads[k].copy_(g.half())
        for tb in tbs:
            check(lib.ngpb_testbed_train_end(tb._h, batch))
        if step % 16 == 0:
            out.append((tbs[0].loss, tbs[0].measured_batch_size, tbs[0].rays_per_batch))
    for tb in tbs:
        tb.close()
    return out


for name, fn in (("single+prefetch", lambda: run_single(True)), ("single", lambda: run_single(False)), ("sharded", run_sharded),
                 ("single+prefetch", lambda: run_single(True)), ("sharded", run_sharded)):
    r = fn()
    print("%-16s" % name, " ".join("%.5f/%d/%d" % x for x in r), flush=True)
