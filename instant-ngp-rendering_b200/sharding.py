"""Host-side sharding arithmetic and collectives of the multi-GPU paths (one process per GPU).

Training shards the rays of a step; rendering shards 256-pixel tiles. The reference is single-GPU
for training (README.md:250-252) and only splits VR views across devices
(src/testbed.cu:5025-5090), so these helpers have no reference counterpart; they mirror what the
kernels do (nerf_train.cu `ray_begin`, nerf_render.cu `kRenderTile`) so tests can check them on CPU.
"""
import numpy as np

RENDER_TILE = 256


def ray_shard(rank, world, n_rays):
    """[begin, end) of the step's global ray indices computed by `rank` (n_rays is a multiple of 256)."""
    if world < 1 or not 0 <= rank < world or 256 % world != 0 or n_rays % 256 != 0:
        raise ValueError("world must divide 256, rank < world, n_rays a multiple of 256")
    per = n_rays // world
    return rank * per, (rank + 1) * per


def tile_shard_pixels(rank, world, begin, end):
    """Row-major pixel indices of [begin, end) traced by `rank`: tiles t with t % world == rank."""
    pix = np.arange(begin, end, dtype=np.int64)
    tile = (pix - begin) // RENDER_TILE
    return pix[tile % world == rank]


def allreduce_counters(counts, loss_sum=None, group=None):
    """Sum of the step counters (and, on the steps that report a loss, the loss sum) over the ranks:
    the first, tiny exchange of a step (Testbed.train_distributed)."""
    import torch.distributed as dist
    dist.all_reduce(counts, group=group)
    if loss_sum is not None:
        dist.all_reduce(loss_sum, group=group)


def allreduce_gradients(gradients, group=None):
    """Sum of the fp16 gradients of all parameters over the ranks (NCCL over NVLink on GPUs)."""
    import torch.distributed as dist
    dist.all_reduce(gradients, group=group)


def allreduce_step(gradients, counts, loss_sum, group=None):
    """Sum of one training step over the ranks: fp16 gradients (all parameters), the three integer
    step counters and the loss sum. torch.distributed picks NCCL (NVLink) for CUDA tensors."""
    import torch.distributed as dist
    dist.all_reduce(gradients, group=group)
    dist.all_reduce(counts, group=group)
    dist.all_reduce(loss_sum, group=group)
