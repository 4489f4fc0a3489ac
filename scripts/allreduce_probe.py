"""Time of the per-step gradient all-reduce (12 865 536 fp16 values = 25.7 MB) and of the 12-byte
counter all-reduce under the current NCCL settings. Launch with torchrun; rank 0 prints."""
import os
import sys

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 12865536
g = torch.randn(n, device="cuda").half()
c = torch.ones(3, dtype=torch.int32, device="cuda")
for t, name in ((g, f"{n} x fp16"), (c, "3 x int32")):
    for _ in range(20):
        dist.all_reduce(t)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(100):
        dist.all_reduce(t)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 100
    if rank == 0:
        tag = " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("NCCL_")) or "defaults"
        bus = t.numel() * t.element_size() * 2 * (world - 1) / world / (ms * 1e-3) / 1e9
        print(f"[{tag}] world {world}: all_reduce({name}) {ms * 1e3:.1f} us, bus bandwidth {bus:.1f} GB/s", flush=True)
    if t is g:
        g.zero_()
dist.barrier()
dist.destroy_process_group()
