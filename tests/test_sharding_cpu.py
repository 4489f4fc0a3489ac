"""Host logic of the multi-GPU paths on CPU: shard arithmetic, and the step all-reduce over a
world_size-2 gloo group (the same helper the NCCL path calls)."""
import importlib
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def sharding(ngp):
    return importlib.import_module("instant-ngp-rendering_b200.sharding")


def test_ray_shards_partition_the_batch(sharding):
    for world in (1, 2, 4, 8):
        for n_rays in (256, 4096, 1 << 18):
            cuts = [sharding.ray_shard(r, world, n_rays) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n_rays
            assert all(a[1] == b[0] for a, b in zip(cuts[:-1], cuts[1:]))
            assert len({e - b for b, e in cuts}) == 1
    with pytest.raises(ValueError):
        sharding.ray_shard(0, 3, 256)
    with pytest.raises(ValueError):
        sharding.ray_shard(2, 2, 256)


def test_tile_shards_partition_the_pixels(sharding):
    for world in (1, 2, 4, 8):
        for begin, end in ((0, 1920 * 1080), (1000, 1000 + 256 * 5 + 17), (5, 5)):
            parts = [sharding.tile_shard_pixels(r, world, begin, end) for r in range(world)]
            allp = np.sort(np.concatenate(parts))
            assert np.array_equal(allp, np.arange(begin, end))
            if end - begin >= 256 * world:
                sizes = [len(p) for p in parts]
                assert max(sizes) - min(sizes) <= 256


def _worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sharding = importlib.import_module("instant-ngp-rendering_b200.sharding")
    g = torch.Generator().manual_seed(rank)
    grads = torch.randn(10007, generator=g).to(torch.float32)  # gloo has no fp16 sum on every build: fp32 stands in
    counts = torch.tensor([100 + rank, 90 + rank, 7 + rank], dtype=torch.int32)
    loss = torch.tensor([0.25 * (rank + 1)], dtype=torch.float32)
    sharding.allreduce_step(grads, counts, loss)
    # the three-phase protocol of Testbed.train_distributed: counters first, gradients later
    counts2 = torch.tensor([5 + rank, 4, 3 * rank], dtype=torch.int32)
    loss2 = torch.tensor([1.5], dtype=torch.float32)
    untouched = torch.tensor([7.0 + rank], dtype=torch.float32)
    sharding.allreduce_counters(counts2, loss2)
    sharding.allreduce_counters(counts2.clone(), None)   # a step that reports no loss
    grads2 = torch.full((33,), float(rank + 1))
    sharding.allreduce_gradients(grads2)
    assert counts2.tolist() == [11, 8, 3] and loss2.item() == 3.0 and untouched.item() == 7.0 + rank
    assert torch.equal(grads2, torch.full((33,), 3.0))
    out.put((rank, grads.numpy().copy(), counts.numpy().copy(), loss.numpy().copy()))
    dist.destroy_process_group()


def test_step_allreduce_gloo_world2(ngp):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in range(2)], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = sum(torch.randn(10007, generator=torch.Generator().manual_seed(r)).numpy() for r in range(2))
    for _, g, c, l in res:
        np.testing.assert_allclose(g, want, rtol=1e-6)
        assert c.tolist() == [201, 181, 15]
        assert abs(float(l[0]) - 0.75) < 1e-7
    # every rank ends with identical buffers: the optimizer then steps identical parameters
    assert np.array_equal(res[0][1], res[1][1])
