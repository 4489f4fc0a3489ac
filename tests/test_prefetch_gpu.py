"""The sample prefetch of Testbed.train (rays of step k+1 marched on a second stream under the
backward pass and the optimizer of step k) must not change what a step computes.

With the parameters frozen (train_network = train_encoding = 0) a training run is a pure function of
the seed: the occupancy grid, the sampled rays, the sample counts and the ray-count control loop are
all integer / deterministic-float work. Two testbeds that differ only in `train_prefetch` must then
report the same (rays_per_batch, samples before compaction, samples after compaction) at every step
- bit-exact - also across the events that force a prefetch to be dropped: occupancy-grid refreshes
(every step at first, then every 2nd ... 16th), a property change, a camera change, an explicit
update_density_grid call and a different batch size.
"""
import importlib

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _run(ngp, prefetch):
    syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
    tb = ngp.Testbed()
    scene, poses, focal = syn.make_dataset(tb, n_images=6, width=64, height=48, aabb_scale=2)
    tb.reload_network_from_json(dict(syn.NERF_BASE_CONFIG))
    tb._set("train_prefetch", prefetch)
    tb._set("train_network", 0)
    tb._set("train_encoding", 0)
    # Few rays at first: a batch that overflows the sample buffer keeps whichever rays won the race
    # for slots (as in the reference), which no two runs repeat.
    tb._set("rays_per_batch", 256)
    trace = []
    params0 = tb.device_buffer("params", torch.float16).clone()

    def step(batch=1 << 18):
        tb.train(batch)
        assert tb.measured_batch_size_before_compaction < 16 * batch, "sample buffer overflowed: the comparison below means nothing"
        trace.append((tb.rays_per_batch, tb.measured_batch_size_before_compaction, tb.measured_batch_size, int(tb._get("training_step"))))

    for _ in range(70):            # prep interval 1 (steps < 32), then 2, 3, 4: prefetched and plain steps interleave
        step()
    tb._set("near_distance", 0.3)  # generator input changes after the prefetch was launched
    for _ in range(3):
        step()
    tb._set("snap_to_pixel_centers", 0)
    step()
    tb.nerf.training.set_camera_extrinsics(2, poses[3], convert_to_ngp=False)
    for _ in range(3):
        step()
    tb.update_density_grid(0.9, 1 << 16, 1 << 16)
    for _ in range(3):
        step()
    step(1 << 17)                  # another batch size: max_samples differs
    for _ in range(4):
        step(1 << 17)
    counters = tb.device_buffer("counters", torch.int32).cpu().numpy().copy()
    bitfield = tb.device_buffer("density_grid_bitfield", torch.uint8).cpu().numpy().copy()
    assert torch.equal(tb.device_buffer("params", torch.float16), params0), "frozen parameters moved"
    tb.close()
    return trace, counters, bitfield


def test_prefetch_does_not_change_the_steps(ngp):
    on, counters_on, bits_on = _run(ngp, 1)
    off, counters_off, bits_off = _run(ngp, 0)
    assert len(on) == len(off) == 85
    for k, (a, b) in enumerate(zip(on, off)):
        assert a == b, (k, a, b)
    # the run did real work and the control loop moved
    assert min(t[2] for t in on) > 0 and len({t[0] for t in on}) >= 3 and len({t[1] for t in on}) > 20
    # counters of the last step (numsteps, numsteps_compacted, rays) and the occupancy bits agree
    assert np.array_equal(counters_on[:3], counters_off[:3])
    assert counters_on[0] == on[-1][1] and counters_on[1] == on[-1][2]
    assert np.array_equal(bits_on, bits_off)


def test_prefetch_with_training(ngp):
    """With the optimizer on, prefetched training still converges like the serial order (the two
    differ only through the atomics' summation order, as two serial runs do)."""
    syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
    losses = {}
    for prefetch in (1, 0):
        tb = ngp.Testbed()
        syn.make_dataset(tb, n_images=8, width=96, height=72, aabb_scale=1)
        tb.reload_network_from_json(dict(syn.NERF_BASE_CONFIG))
        tb._set("train_prefetch", prefetch)
        for _ in range(161):
            tb.train(1 << 14)
        losses[prefetch] = tb.loss
        adopted = tb._get("prefetches_adopted")
        assert adopted >= 80 if prefetch else adopted == 0, adopted
        tb.close()
    assert losses[1] > 0 and losses[0] > 0
    assert losses[1] < 2.0 * losses[0] and losses[0] < 2.0 * losses[1], losses
