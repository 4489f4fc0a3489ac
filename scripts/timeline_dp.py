"""Kernel timeline of sharded NeRF training steps on rank 0 (diagnosis only; launch with torchrun).
Usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/timeline_dp.py [pretrain] [traced]"""
import importlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from torch.profiler import ProfilerActivity, profile

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
pre = int(sys.argv[1]) if len(sys.argv) > 1 else 200
traced = int(sys.argv[2]) if len(sys.argv) > 2 else 3
batch = (1 << 18) * world
tb = ngp.Testbed()
tb.use_current_stream()
syn.make_dataset(tb, n_images=20, width=960, height=540, aabb_scale=4, amp_scale=8.0)
tb.reload_network_from_json(syn.NERF_BASE_CONFIG)
tb.set_shard(rank, world)
for _ in range(pre):
    tb.train_distributed(batch)
while tb.training_step % 16 != 1:
    tb.train_distributed(batch)
torch.cuda.synchronize()
dist.barrier()
t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t0.record()
for _ in range(15):
    tb.train_distributed(batch)
t1.record()
torch.cuda.synchronize()
if rank == 0:
    print(f"world {world}: {t0.elapsed_time(t1) / 15:.4f} ms/step, rays_per_batch {tb.rays_per_batch}")
while tb.training_step % 16 != 1:
    tb.train_distributed(batch)
torch.cuda.synchronize()
dist.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(traced):
        tb.train_distributed(batch)
    torch.cuda.synchronize()
if rank == 0:
    out = os.path.join(ROOT, "gpurun_out", "timeline_dp.json")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    prof.export_chrome_trace(out)
    ev = json.load(open(out))["traceEvents"]
    ks = sorted((e for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset") and "ts" in e), key=lambda e: e["ts"])
    z = ks[0]["ts"] if ks else 0
    for e in ks:
        a = e.get("args", {})
        print(f"{e['ts'] - z:10.1f} us  +{e['dur']:8.1f} us  stream {a.get('stream', '?'):>3}  grid {str(a.get('grid', '')):>16} {e['name'][:60]}")
    cpu = sorted((e for e in ev if e.get("cat") in ("cuda_runtime", "cuda_driver") and "ts" in e), key=lambda e: e["ts"])
    print("host API calls:", len(cpu), "spanning", (cpu[-1]["ts"] - cpu[0]["ts"]) if cpu else 0, "us")
dist.barrier()
dist.destroy_process_group()
