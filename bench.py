#!/usr/bin/env python
"""bench.py - throughput of the ngp-b200 hot path on N GPUs of one node.

    python bench.py --gpus N --steps K --warmup W [--workload nerf|grid] [--impl reference]

Prints ONE JSON line (rank 0). DESIGN.md "Measurement" says what each key means.

Workloads (BASELINE.json configs):
  nerf   (default) configs[2] + configs[3]: NeRF training with configs/nerf/base.json on a synthetic
         dataset of the fox shape (50 views 1920x1080 RGBA8, aabb_scale 4), 2^18 samples per batch,
         through Testbed.train(); headline metric "fox NeRF train samples/s" (compacted samples / s).
         The same line carries the novel-view render of the trained model at 1920x1080 ("render",
         Mrays/s) and the hash-grid micro-benchmark ("grid", Gpts/s).
         N > 1: rays of every step are sharded over the ranks and the gradients all-reduced over
         NCCL; the render shards pixel tiles with no communication.
         It also carries configs[1] ("image": neural image of the albert shape, samples/s).
  grid   configs[0]: hash-grid encoding forward+backward alone, 3-D L16 F2 T=2^19, 2^20 points.
         Replicas only at N > 1 (no exchange).
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "baseline"))
import configs as ref_configs  # noqa: E402  the reference's shipped configs restated; imports nothing of the product

def use_host_cores(oracle):
    """torchrun pins OMP_NUM_THREADS=1 in its workers; the timed CPU legs (the oracle's OpenMP loops) are meant to use the
    host's cores. Set on the oracle library only: raising the variable for the whole process made every rank's torch
    thread pool spin on all cores and doubled the 8-GPU render time."""
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    oracle.set_num_threads(n)


FOX_DIR = os.path.join(ROOT, "data_stage", "fox")   # the reference's data/nerf/fox, staged (git-ignored); absent -> synthetic fox-shaped set

GRID_ENC = {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 19,
            "base_resolution": 16, "per_level_scale": 1.3819}
GRID_N = 1 << 20
GRID_BYTES_PER_POINT_FWD = 12 + 512 + 64    # xyz + 16 levels x 8 corners x 4 B + 32 fp16 out
GRID_BYTES_PER_POINT_BWD = 12 + 64 + 512    # xyz + dL/dy + scatter RMW counted once

# fox shape (data/nerf/fox: 50 views 1080 x 1920, aabb_scale 4) and configs/nerf/base.json
FOX = {"n_images": 50, "width": 1920, "height": 1080, "aabb_scale": 4, "batch": 1 << 18, "amp_scale": 8.0,
       "pretrain_steps": int(os.environ.get("NGPB_BENCH_PRETRAIN", "512"))}  # the override exists for short ncu captures
NERF_MLP_FLOP_TRAIN = 61440      # per trained sample, fwd+bwd
NERF_MLP_FLOP_INFER = 20480      # per marched sample
NERF_GRID_BYTES_TRAIN = 1176     # per trained sample (encoder fwd+bwd)
NERF_GRID_BYTES_INFER = 588      # per marched sample


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d.get("bf16_tflops"), "bf16_tflops_sustained": d.get("bf16_tflops_sustained"), "source": "measured"}
    # fallback stated in /opt/skills/guides/B200_PROFILING.md
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """Samples nvidia-smi clocks and throttle reasons while the timed region runs."""

    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._t = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([s.strip() for s in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.05)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                mx = max(mx, float(s[1]))
                for nme, v in zip(names, s[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup():
    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        # stdout carries ONE JSON line: whatever NCCL logs (its version banner under NCCL_DEBUG=VERSION / INFO) goes to stderr
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        if not dist.is_initialized():
            dist.init_process_group(backend="nccl" if torch.cuda.is_available() else "gloo")
    return rank, world, local


def barrier(world):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if torch.cuda.is_available():
        torch.cuda.synchronize()


def reduce_over_ranks(value, world, op="max"):
    import torch
    if world == 1:
        return value
    import torch.distributed as dist
    t = torch.tensor([value], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return float(t.item())


# ------------------------------------------------------------------------------------------
# CPU baselines: the oracle (a CPU restatement of the reference's kernels; the reference itself has
# no CPU implementation of this path) timed on the host cores. Used by the
# `cpu_baseline` object of the GPU arm and by `--impl reference`; never by the product path.
def cpu_grid_baseline(budget_points):
    import numpy as np
    import oracle_binding as oracle
    use_host_cores(oracle)
    cfg = oracle.grid_cfg(3, GRID_ENC)
    n_entries = int(oracle.grid_offsets(cfg)[-1])
    rng = np.random.default_rng(1337)
    x = rng.random((budget_points, 3), dtype=np.float32)
    params = (rng.standard_normal(n_entries * 2) * 0.1).astype(np.float16)
    dy = rng.standard_normal((budget_points, 32)).astype(np.float16)
    t0 = time.perf_counter()
    oracle.grid_forward(cfg, x, params)
    oracle.grid_backward(cfg, x, dy)
    dt = time.perf_counter() - t0
    return budget_points / dt / 1e9, dt, oracle.num_threads()


class CpuNerf:
    """One NeRF training step of the reference's algorithm on the host, composed from the oracle:
    generate_training_samples_nerf -> NerfNetwork inference -> compute_loss_kernel_train_nerf ->
    NerfNetwork forward+backward on the compacted batch -> Adam. Inputs have the fox shape
    (50 x 1080 x 1920 RGBA8, aabb_scale 4, base.json network); the step is run on a bounded number
    of rays so that one step costs seconds, and the metric is compacted samples per second."""

    def __init__(self, n_rays):
        import numpy as np
        import oracle_binding as oracle
        self.np, self.oracle = np, oracle
        syn = ref_configs   # poses and configs only: the CPU arm must not load the product library
        rng = np.random.default_rng(1337)
        n_images, w, h, aabb = FOX["n_images"], FOX["width"], FOX["height"], FOX["aabb_scale"]
        self.data = "synthetic"
        if os.path.isdir(FOX_DIR) and not os.environ.get("NGPB_BENCH_SYNTHETIC"):
            # the reference's own fox scene (staged), read by the loader's host logic loaded by PATH: the package itself
            # would map libngp_b200.so, which the CPU arm must not do
            import importlib.util
            spec = importlib.util.spec_from_file_location("ngpb_loader_host", os.path.join(ROOT, "instant-ngp-rendering_b200", "loader.py"))
            loader = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(loader)
            ds = loader.describe_dataset(FOX_DIR)
            self.images = []
            for fr in ds["frames"]:
                px = loader.load_frame_pixels(fr, ds)
                fx, fy = loader.finish_frame(fr, px.shape[1], px.shape[0])
                self.images.append(oracle.nerf_image(px, fr["xform"], focal=(fx, fy), principal=tuple(fr["principal"]),
                                                     lens=fr["lens_params"][:4] if fr["lens_mode"] == 1 else None))
            n_images, aabb = len(self.images), int(ds["aabb_scale"])
            self.data = "fox"
        else:
            poses = syn.orbit_poses(n_images, radius=1.4)
            # procedural host images: one random RGBA8 tile repeated (the kernels only look pixels up)
            tile = rng.integers(0, 256, size=(h, w, 4), dtype=np.uint8)
            tile[..., 3] = 255
            self.images = [oracle.nerf_image(tile, poses[i], focal=1.1 * w) for i in range(n_images)]
        self.max_cascade = int(np.log2(aabb))
        # occupancy of a partly trained model: 8x8x8 bricks (contiguous Morton runs) in every cascade
        grid = np.zeros((self.max_cascade + 1, 128 ** 3), dtype=np.float32)
        for c in range(self.max_cascade + 1):
            for s in rng.integers(0, 128 ** 3 // 512, size=400) * 512:
                grid[c, s:s + 512] = 0.5
        self.bitfield, self.mean = oracle.nerf_bitfield(self.max_cascade, grid.reshape(-1))
        enc = dict(syn.NERF_BASE_CONFIG["encoding"])
        enc["per_level_scale"] = float(np.exp(np.log(2048.0 * aabb / enc["base_resolution"]) / (enc["n_levels"] - 1)))
        self.gcfg = oracle.grid_cfg(3, enc)
        n_grid = int(oracle.grid_offsets(self.gcfg)[-1]) * enc["n_features_per_level"]
        st = oracle.Pcg32State(1337)
        wd = oracle.mlp_init(st, 32, 1)
        wr = oracle.mlp_init(st, 32, 2)
        wg = oracle.random_uniform_from(st, n_grid, -1e-4, 1e-4)
        self.params32 = np.concatenate([wd, wr, wg]).astype(np.float32)
        self.n_mlp = len(wd) + len(wr)
        self.m = np.zeros_like(self.params32)
        self.v = np.zeros_like(self.params32)
        self.step_count = 0
        cfg = oracle.NerfConfig()
        half = 0.5 * aabb
        for k in range(3):
            cfg.aabb_min[k], cfg.aabb_max[k], cfg.background_color[k] = 0.5 - half, 0.5 + half, 0.0
        cfg.n_rays, cfg.n_rays_total = n_rays, 0
        cfg.max_samples, cfg.max_samples_compacted = 1 << 22, 1 << 22
        cfg.n_training_images, cfg.max_mip = n_images, self.max_cascade
        cfg.cone_angle_constant = 1.0 / 256.0
        cfg.near_distance, cfg.loss_scale = 0.1, 128.0
        cfg.snap_to_pixel_centers, cfg.random_bg_color, cfg.linear_colors, cfg.color_space = 1, 1, 0, 0
        cfg.loss_type, cfg.rgb_activation, cfg.density_activation = 4, 2, 3
        self.cfg = cfg
        self.rng = oracle.Pcg32State(1337)

    def step(self):
        np, oracle, cfg = self.np, self.oracle, self.cfg
        params16 = self.params32.astype(np.float16)
        numsteps, rays, coords, total = oracle.nerf_generate_samples(cfg, self.images, self.bitfield, self.rng, cfg.max_samples)
        ray_idx = np.nonzero(numsteps)[0].astype(np.uint32)
        base = np.concatenate([[0], np.cumsum(numsteps.astype(np.int64))])[:-1]
        ns_in = np.stack([numsteps[ray_idx], base[ray_idx].astype(np.uint32)], axis=1).astype(np.uint32)
        fwd = oracle.nerf_forward(self.gcfg, params16, coords, 1, 2)
        rgbsigma = np.ascontiguousarray(fwd["out"][:, :4])
        comp, loss, dloss = oracle.nerf_compute_loss(cfg, self.images, self.mean, ray_idx, ns_in, rays[ray_idx], coords, rgbsigma, self.rng)
        sel = np.concatenate([np.arange(b, b + c) for b, c in zip(ns_in[:, 1].astype(np.int64), comp.astype(np.int64))]) if len(comp) else np.zeros(0, dtype=np.int64)
        cc = np.ascontiguousarray(coords[sel])
        dy = np.ascontiguousarray(dloss[sel]).astype(np.float16)
        fwd2 = oracle.nerf_forward(self.gcfg, params16, cc, 1, 2)
        dWd, dWr, gg, _ = oracle.nerf_backward(self.gcfg, fwd2, cc, dy, 1, 2)
        grad = np.concatenate([dWd.reshape(-1), dWr.reshape(-1), gg.reshape(-1)]).astype(np.float32) / np.float32(cfg.loss_scale)
        # Adam (tcnn adam.h:48-119): beta 0.9 / 0.99, eps 1e-15, lr 1e-2, L2 1e-6 on the MLP weights
        self.step_count += 1
        grad[:self.n_mlp] += np.float32(1e-6) * self.params32[:self.n_mlp]
        self.m = np.float32(0.9) * self.m + np.float32(0.1) * grad
        self.v = np.float32(0.99) * self.v + np.float32(0.01) * grad * grad
        lr = 1e-2 * np.sqrt(1 - 0.99 ** self.step_count) / (1 - 0.9 ** self.step_count)
        self.params32 -= np.float32(lr) * self.m / (np.sqrt(self.v) + np.float32(1e-15))
        self.rng.advance(1 << 32)
        return int(comp.sum()), int(total), float(loss.sum())


def cpu_nerf_baseline(n_rays, steps, warmup=0, budget_s=1e9):
    """Runs up to `steps` oracle training steps (stops early once `budget_s` of timed work is spent)."""
    import oracle_binding as oracle
    use_host_cores(oracle)
    c = CpuNerf(n_rays)
    for _ in range(warmup):
        c.step()
    samples, marched, dt, done = 0, 0, 0.0, 0
    for _ in range(steps):
        t0 = time.perf_counter()
        s, m, _ = c.step()
        dt += time.perf_counter() - t0
        samples += s
        marched += m
        done += 1
        if dt > budget_s:
            break
    return {"samples_per_s": samples / dt, "seconds": dt, "steps_done": done, "samples": samples, "marched": marched, "cores": oracle.num_threads(), "data": c.data,
            "sample": f"{done} oracle training step(s) of {n_rays} rays ({samples // max(done, 1)} compacted / {marched // max(done, 1)} marched samples per step, "
                      f"{dt:.1f} s) on the {'reference fox scene (data/nerf/fox, staged)' if c.data == 'fox' else 'fox-shaped synthetic inputs'}"}


# ------------------------------------------------------------------------------------------
# GPU arm
def grid_measure(ngp, steps, warmup, rank, world, local):
    """configs[0]: hash-grid forward+backward alone; returns the fields of a bench line."""
    import ctypes
    import numpy as np
    import torch
    peaks = load_peaks()
    dev = torch.device("cuda", local)
    n = GRID_N
    e = ngp.Encoding(3, GRID_ENC, device=dev)
    g = torch.Generator(device=dev).manual_seed(1337 + rank)
    x = torch.rand((n, 3), generator=g, device=dev)
    params = (torch.randn(e.n_params, generator=g, device=dev) * 0.1).half()
    dy = (torch.randn((n, 32), generator=g, device=dev) * (128.0 / n)).half()
    e.set_params(params)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    lib, check = ngp.lib, ngp._lib.check
    ev = lambda: torch.cuda.Event(enable_timing=True)
    out = torch.empty((n, 32), dtype=torch.float16, device=dev)
    grad = torch.empty(e.n_params, dtype=torch.float16, device=dev)
    sp = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def fwd_bwd(record=None):
        ctx = ctypes.c_void_p()
        if record: record[0].record()
        check(lib.ngpb_module_forward(e._h, sp, n, x.data_ptr(), out.data_ptr(), params.data_ptr(), 0, ctypes.byref(ctx)))
        if record: record[1].record()
        check(lib.ngpb_module_backward(e._h, sp, ctx, n, None, dy.data_ptr(), grad.data_ptr(), x.data_ptr(), out.data_ptr(), params.data_ptr()))
        if record: record[2].record()
        return ctx

    for _ in range(max(warmup, 3)):
        ctx = fwd_bwd()
        torch.cuda.synchronize()
        lib.ngpb_context_free(ctx)
    fwd_ms, bwd_ms, step_ms = [], [], []
    barrier(world)
    for _ in range(steps):
        flush.fill_(1)
        rec = (ev(), ev(), ev())
        ctx = fwd_bwd(rec)
        torch.cuda.synchronize()
        lib.ngpb_context_free(ctx)
        fwd_ms.append(rec[0].elapsed_time(rec[1]))
        bwd_ms.append(rec[1].elapsed_time(rec[2]))
        step_ms.append(rec[0].elapsed_time(rec[2]))
    barrier(world)
    total_ms = reduce_over_ranks(sum(step_ms), world)
    value = n * world * steps / (total_ms * 1e-3) / 1e9

    # end to end through the module API with pinned host buffers, copies inside the timed region
    xh, dyh = x.cpu().pin_memory(), dy.cpu().pin_memory()
    outh = torch.empty((n, 32), dtype=torch.float16).pin_memory()
    gradh = torch.empty(e.n_params, dtype=torch.float16).pin_memory()

    def e2e_step():
        xd = xh.to(dev, non_blocking=True)
        dyd = dyh.to(dev, non_blocking=True)
        o, c = e.forward(xd)
        _, gr = e.backward(c, xd, o, dyd)
        outh.copy_(o, non_blocking=True)
        gradh.copy_(gr, non_blocking=True)
    for _ in range(3):
        e2e_step()
    barrier(world)
    t0, t1 = ev(), ev()
    t0.record()
    for _ in range(steps):
        e2e_step()
    t1.record()
    torch.cuda.synchronize()
    e2e_ms = reduce_over_ranks(t0.elapsed_time(t1), world)
    # gather / scatter rooflines measured in the same run: random 4-byte loads and random
    # red.f16x2 over a 32 MiB L2-resident table (the encoder's table is 24.4 MB), 2^20 x 128 accesses
    table = torch.zeros(1 << 23, dtype=torch.int32, device=dev)
    sink = torch.zeros(4, dtype=torch.int32, device=dev)
    probe = {}
    for name, call in (("gather", lambda: lib.ngpb_probe_gather(sp, ctypes.c_void_p(table.data_ptr()), 23, 4, 1 << 20, 128, ctypes.c_void_p(sink.data_ptr()))),
                       ("scatter", lambda: lib.ngpb_probe_scatter(sp, ctypes.c_void_p(table.data_ptr()), 23, 1 << 20, 128))):
        for _ in range(2):
            check(call())
        t0, t1 = ev(), ev()
        t0.record()
        for _ in range(5):
            check(call())
        t1.record()
        torch.cuda.synchronize()
        probe[name] = 5 * (1 << 20) * 128 / (t0.elapsed_time(t1) * 1e-3) / 1e9   # G accesses / s
    # L2 -> SM streaming bandwidth, same run: 64 coalesced passes over the same 32 MiB (ld.global.cg, 16 B per lane)
    l2_passes = 64
    for _ in range(2):
        check(lib.ngpb_probe_l2_stream(sp, ctypes.c_void_p(table.data_ptr()), 32 << 20, l2_passes, 148 * 8, ctypes.c_void_p(sink.data_ptr())))
    t0, t1 = ev(), ev()
    t0.record()
    check(lib.ngpb_probe_l2_stream(sp, ctypes.c_void_p(table.data_ptr()), 32 << 20, l2_passes, 148 * 8, ctypes.c_void_p(sink.data_ptr())))
    t1.record()
    torch.cuda.synchronize()
    l2_peak = (32 << 20) * l2_passes / (t0.elapsed_time(t1) * 1e-3) / 1e9   # GB/s
    f_ms, b_ms = float(np.mean(fwd_ms)), float(np.mean(bwd_ms))
    sectors = load_grid_sectors()
    fwd_rate = n * 128 / (f_ms * 1e-3) / 1e9
    bwd_rate = n * 128 / (b_ms * 1e-3) / 1e9
    back = b_ms >= f_ms
    dom = "grid_backward_kernel" if back else "grid_forward_kernel"
    achieved = n * (GRID_BYTES_PER_POINT_BWD if back else GRID_BYTES_PER_POINT_FWD) / (max(f_ms, b_ms) * 1e-3) / 1e9
    return {
        "metric": "hash-enc fwd+bwd Gpts/s", "value": value, "unit": "Gpts/s", "ms_per_step": total_ms / steps,
        "config": {"workload": "grid_L16F2_T19_2^20pts", "encoding": GRID_ENC, "points_per_step_per_gpu": n,
                   "l2": "flushed between timed steps (256 MiB write)", "parallelism": "replicas only" if world > 1 else "1 GPU"},
        "kernels_ms": {"grid_forward_kernel": f_ms, "memset+grid_backward_kernel": b_ms},
        "gather_roofline": {"what": "16 levels x 8 corners = 128 4-byte accesses per point against the rate of FULLY RANDOM accesses (probe kernels, same run, 32 MiB table); "
                                    "the encoder beats the random rate forward because the two x-neighbours of a corner pair share a sector and the coarse levels hit L1",
                            "forward": {"achieved": fwd_rate, "peak": probe["gather"], "unit": "G loads/s", "frac": fwd_rate / probe["gather"]},
                            "backward": {"achieved": bwd_rate, "peak": probe["scatter"], "unit": "G red.f16x2/s", "frac": bwd_rate / probe["scatter"]},
                            "fwd_bwd": {"ms": f_ms + b_ms, "ms_at_probe_rates": (n * 128 / probe["gather"] + n * 128 / probe["scatter"]) / 1e9 * 1e3,
                                        "frac": (n * 128 / probe["gather"] + n * 128 / probe["scatter"]) / 1e9 * 1e3 / (f_ms + b_ms)}},
        "roofline": grid_sector_roofline(sectors, f_ms, b_ms, l2_peak, n),
        "hbm_floor": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                      "traffic": load_traffic(dom), "peak_source": peaks["source"],
                      "note": "algorithmic 588 B/point per direction; the 24.4 MB table is L2-resident, so the HBM copy peak is a floor for the gather roofline, not a ceiling"},
        "e2e": {"value": n * world * steps / (e2e_ms * 1e-3) / 1e9, "unit": "Gpts/s", "h2d_bytes_per_step": int(xh.numel() * 4 + dyh.numel() * 2),
                "d2h_bytes_per_step": int(outh.numel() * 2 + gradh.numel() * 2)},
        "gpu_launches": 3 * steps,   # grid_forward_kernel, grid_backward_coarse_kernel, grid_backward_kernel
    }


def load_grid_sectors():
    """L2 sectors per launch of the config-1 kernels from the committed ncu capture (profiles/grid_sectors.json), or None."""
    tp = os.path.join(ROOT, "profiles", "grid_sectors.json")
    try:
        return json.load(open(tp))
    except Exception:
        return None


def grid_sector_roofline(sectors, f_ms, b_ms, l2_peak, n):
    """The hash grid's table lives in L2, so the roof is L2 bandwidth: sectors the kernels move through L2 (lts__t_sectors.sum of the
    committed ncu capture of this workload, x 32 B) over the kernel time measured live, against the streaming L2 -> SM rate the
    probe measured in this run."""
    if not sectors:
        return {"bound": "l2", "achieved": None, "peak": l2_peak, "unit": "GB/s", "frac": None, "traffic": None, "note": "profiles/grid_sectors.json missing"}
    sf = float(sectors["grid_forward_kernel"]["lts_sectors"])
    sb = float(sectors["grid_backward_kernel"]["lts_sectors"]) + float(sectors.get("grid_backward_coarse_kernel", {}).get("lts_sectors", 0))
    scale = n / float(sectors.get("points", n))
    fwd = sf * scale * 32 / (f_ms * 1e-3) / 1e9
    bwd = sb * scale * 32 / (b_ms * 1e-3) / 1e9
    both = (sf + sb) * scale * 32 / ((f_ms + b_ms) * 1e-3) / 1e9
    dram = sum(float(v.get("dram_bytes", 0)) for v in sectors.values() if isinstance(v, dict))
    return {"bound": "l2", "kernel": "grid_forward_kernel + grid_backward_coarse_kernel + grid_backward_kernel", "achieved": both, "peak": l2_peak, "unit": "GB/s",
            "frac": both / l2_peak, "traffic": dram * scale, "peak_source": "measured in this run (ngpb_probe_l2_stream: 64 passes over 32 MiB, ld.global.cg)",
            "forward": {"achieved": fwd, "frac": fwd / l2_peak, "sectors": sf * scale}, "backward": {"achieved": bwd, "frac": bwd / l2_peak, "sectors": sb * scale},
            "note": "sector-based: lts__t_sectors.sum per launch (profiles/grid_sectors.json, ncu capture of this workload) x 32 B / live kernel time; "
                    "traffic = DRAM bytes of the same capture (the table is L2-resident: mostly the flushed first touch, the inputs and the outputs)"}


IMAGE_CONFIG = {  # configs/image/base.json
    "loss": {"otype": "L2"},
    "optimizer": {"otype": "ExponentialDecay", "decay_start": 20000, "decay_interval": 10000, "decay_base": 0.33,
                  "nested": {"otype": "Adam", "learning_rate": 1e-2, "beta1": 0.9, "beta2": 0.99, "epsilon": 1e-15, "l2_reg": 1e-6}},
    "encoding": {"otype": "HashGrid", "n_levels": 16, "n_features_per_level": 2, "log2_hashmap_size": 24, "base_resolution": 16},
    "network": {"otype": "FullyFusedMLP", "activation": "ReLU", "output_activation": "None", "n_neurons": 64, "n_hidden_layers": 2},
}


def image_measure(ngp, steps, warmup, rank, world, local):
    """configs[1]: neural image of the albert shape (1024 x 1024 RGBA float), base.json, 2^18 samples per
    step through Testbed.train; replicas only at N > 1."""
    import numpy as np
    import torch
    dev = torch.device("cuda", local)
    w = h = 1024
    ys, xs = torch.meshgrid(torch.arange(h, device=dev), torch.arange(w, device=dev), indexing="ij")
    g = torch.Generator(device=dev).manual_seed(7)
    img = torch.stack([0.5 + 0.5 * torch.sin(xs / 37.0) * torch.cos(ys / 23.0), ((xs // 32 + ys // 32) % 2) * 0.8 + 0.1,
                       (xs / w + 0.05 * torch.randn((h, w), generator=g, device=dev)).clamp(0, 1), torch.ones((h, w), device=dev)], dim=-1).float().contiguous()
    tb = ngp.Testbed(device=dev)
    tb.use_current_stream()
    albert = os.path.join(ROOT, "data_stage", "albert.exr")   # the reference's data/image/albert.exr, staged
    if os.path.exists(albert) and not os.environ.get("NGPB_BENCH_SYNTHETIC"):
        loader = importlib.import_module("instant-ngp-rendering_b200.loader")
        img = torch.from_numpy(loader.load_image_file(albert)).to(dev).float().contiguous()
        h, w = int(img.shape[0]), int(img.shape[1])
        name = "albert.exr"
    else:
        name = "synthetic 1024x1024"
    tb.set_image(img)
    tb.reload_network_from_json(IMAGE_CONFIG)
    batch = 1 << 18
    for _ in range(max(warmup, 3) + 100):
        tb.train(batch)
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        tb.train(batch)
    e1.record()
    torch.cuda.synchronize()
    ms = reduce_over_ranks(e0.elapsed_time(e1), world)
    got = tb.render_image(w, h)
    mse = float(((got[..., :3].clamp(0, 1) - img[..., :3]) ** 2).mean())
    tb.close()
    peaks = load_peaks()
    pk = peaks["bf16_tflops_sustained"] or peaks["bf16_tflops"]
    tf = batch * 43008.0 / (ms / steps * 1e-3) / 1e12     # 32 -> 64 -> 64 -> 16 forward + backward, per GPU
    return {"metric": "neural image train samples/s", "value": batch * world * steps / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms / steps,
            "config": {"workload": "image_%s_base.json_batch2^18" % name, "parallelism": "replicas only" if world > 1 else "1 GPU"},
            "roofline": {"bound": "tensor", "achieved": tf, "peak": pk, "unit": "TFLOP/s", "frac": tf / pk, "traffic": None,
                         "note": "whole step (positions, 2-D grid fwd+bwd, MLP fwd+bwd, loss, Adam) against the MLP's 43 008 FLOP per sample"},
            "mse_linear_after": mse, "steps_trained": steps + max(warmup, 3) + 100}


def load_traffic(kernel):
    """dram bytes per launch of `kernel` from the committed ncu --set full capture (profiles/traffic.json), or None."""
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        return json.load(open(tp)).get(kernel)
    return None


# kernels launched by one Testbed.train step (profiles/r01_nerf_launches.csv lists them): sample
# generation, encoder + fused MLP (inference), loss/compaction + roll-over, encoder, fused MLP
# fwd+bwd (+ partial-sum reduction), grid backward, optimizer; the occupancy update adds 9 every
# 16th step and the loss reduction 1.
NERF_LAUNCHES_PER_STEP = 11
# kernels behind one mark of the library's phase timer (csrc: nerf_train.cu, network.cu, nerf_loss.cu, optimizer.cu, testbed_train.cu);
# gpu_launches = sum over the phases of (marks counted by the timer in the profiled steps) x (kernels per mark), scaled to K steps
KERNELS_PER_PHASE = {"dp.reduce_scatter": 1, "dp.all_gather": 1, "optimizer.ema_of_other_shards": 2, "generate_training_samples": 1, "infer.grid_forward": 1, "infer.fused_mlp": 1, "infer.fused_grid_mlp": 1, "loss_and_compact": 2,
                     "train.grid_forward": 1, "train.fused_mlp_fwd_bwd": 2, "train.fused_grid_mlp_fwd_bwd_scatter": 2, "train.grid_backward": 1,
                     "optimizer": 1, "density_grid_update": 9}


def count_launches(prof, profiled_steps, steps):
    n = sum(cnt * KERNELS_PER_PHASE.get(name, 1) for name, (_, cnt) in prof.items())
    return int(round(n * steps / max(profiled_steps, 1)))


def gpu_baseline(timeout_s=240):
    """The reference's OWN CUDA build (baseline/_ref, built by baseline/build_ref.sh from the unmodified sources, native
    sm_100 SASS) on this GPU, through its own Python binding / C++ API, in child processes: fox training at 2^18 samples
    per step (steps 300..1300 timed), the 1920x1080 render of the trained model through render(), albert at 2^18, and
    its kernels on the benchmark shapes (baseline/harness.cu). None when the build or the staged data is absent."""
    ref = os.path.join(ROOT, "baseline", "_ref")
    if not (os.path.exists(os.path.join(ref, "ref_harness")) and any(f.startswith("pyngp") and f.endswith(".so") for f in os.listdir(ref))):
        return {"unavailable": "baseline/_ref is not built here (baseline/build_ref.sh needs /root/reference)"}
    out = {"build": "unmodified reference sources, nvcc 12.9 -gencode arch=compute_100,code=sm_100 (tcnn's arch cap at 90 lifted), --use_fast_math, headless",
           "how": "child processes; CUDA events / device-synchronised wall clock inside them (baseline/run_pyngp.py, baseline/dump_ref_gpu.py)"}
    tmp = os.path.join(ROOT, "gpurun_out")
    os.makedirs(tmp, exist_ok=True)
    env = dict(os.environ)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR", "MASTER_PORT"):
        env.pop(k, None)
    try:
        if os.path.isdir(FOX_DIR):
            what = "fox,albert" if os.path.exists(os.path.join(ROOT, "data_stage", "albert.exr")) else "fox"
            jp = os.path.join(tmp, "bench_ref_pyngp.json")
            subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "run_pyngp.py"), "--impl", "reference", "--what", what, "--steps", "300,1300",
                            "--eval", "0", "--frames", "5", "--out", jp], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=timeout_s, env=env, check=True)
            r = json.load(open(jp))
            last = r["fox"]["checkpoints"][-1]
            out["train"] = {"workload": "data/nerf/fox, configs/nerf/base.json, batch 2^18, steps 300..1300", "ms_per_step": last["ms_per_step"],
                            "samples_per_s": r["fox"]["samples_per_s"], "note": "credited with the full 2^18 batch per step (its binding does not expose the compacted count)"}
            out["render"] = {"workload": "trained fox, training view 0, 1920x1080, 1 spp, through render() incl. the copy to the host", "ms_per_frame": r["fox"]["render"]["ms_per_frame_e2e"],
                             "mrays_per_s": r["fox"]["render"]["mrays_per_s_e2e"]}
            if "albert" in r:
                out["image"] = {"workload": "data/image/albert.exr, configs/image/base.json, batch 2^18, steps 1000..2000", "ms_per_step": r["albert"]["checkpoints"][-1]["ms_per_step"],
                                "samples_per_s": r["albert"]["checkpoints"][-1]["samples_per_s"]}
        else:
            out["train"] = {"unavailable": "data_stage/fox is not staged on this box"}
        jp = os.path.join(tmp, "ref_gpu_timings.json")
        subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "dump_ref_gpu.py"), "--timings-only"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=timeout_s, env=env, check=True)
        t = json.load(open(jp))
        out["grid"] = {"workload": "L16 F2 T=2^19, 2^20 uniform points, L2 flushed per iteration", "fwd_ms": t["grid_L16F2"]["fwd_ms"], "bwd_ms": t["grid_L16F2"]["bwd_ms"],
                       "gpts_per_s": t["grid_L16F2"]["gpts_per_s_fwd_bwd"], "L8F4_gpts_per_s": t["grid_L8F4"]["gpts_per_s_fwd_bwd"]}
        out["kernels"] = {"nerf_network_2^18": t["nerfnet_2e18"], "fully_fused_mlp_2^18": t["mlp_32_64x2_3_2e18"], "adam_12.9M": t["adam_12.9M"]}
    except Exception as e:  # the baseline is a report, never a reason for the bench to fail
        out["error"] = "%s: %s" % (type(e).__name__, str(e)[:200])
    return out


def nerf_phase_costs(prof, n_steps, compacted, marched, n_params, peaks):
    """Per-kernel-group time, algorithmic work and roofline fraction from the library's phase timer."""
    work = {  # name -> (bound, algorithmic units per step, unit)
        "infer.grid_forward": ("hbm", marched * float(NERF_GRID_BYTES_INFER), "B"),
        "infer.fused_mlp": ("tensor", marched * float(NERF_MLP_FLOP_INFER), "FLOP"),
        "loss_and_compact": ("hbm", marched * 12.0 + compacted * 64.0, "B"),
        "train.grid_forward": ("hbm", compacted * float(NERF_GRID_BYTES_INFER), "B"),
        "train.fused_mlp_fwd_bwd": ("tensor", compacted * float(NERF_MLP_FLOP_TRAIN), "FLOP"),
        "train.grid_backward": ("hbm", compacted * float(NERF_GRID_BYTES_INFER) + n_params * 2.0, "B"),
        "optimizer": ("hbm", n_params * 22.0, "B"),
    }
    out = {}
    for name, (ms, cnt) in prof.items():
        per_step = ms / n_steps
        entry = {"ms_per_step": per_step}
        if name in work and per_step > 0:
            bound, units, _ = work[name]
            if bound == "hbm":
                a = units / (per_step * 1e-3) / 1e9
                entry.update({"bound": "hbm", "achieved": a, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": a / peaks["hbm_gbs"]})
            else:
                a = units / (per_step * 1e-3) / 1e12
                pk = peaks["bf16_tflops_sustained"] or peaks["bf16_tflops"]
                entry.update({"bound": "tensor", "achieved": a, "peak": pk, "unit": "TFLOP/s", "frac": a / pk})
        if name == "generate_training_samples" and per_step > 0:
            # a serial loop per ray (about 270 iterations of ~320 instructions for ~50 samples): bound by
            # dependent-issue latency, not by bytes or flops - reported as a rate, outside the rooflines
            entry.update({"bound": "issue latency (serial per-ray march)", "marched_samples_per_s": marched / (per_step * 1e-3),
                          "output_GBps": marched * 28.0 / (per_step * 1e-3) / 1e9,
                          "note": "timed back to back here; in the timed steps above it runs on a second stream under the previous step's backward pass and optimizer"})
        out[name] = entry
    return out


def run_nerf(args):
    import numpy as np
    import torch
    ngp = importlib.import_module("instant-ngp-rendering_b200")
    syn = importlib.import_module("instant-ngp-rendering_b200.synthetic")
    tbm = importlib.import_module("instant-ngp-rendering_b200.testbed")
    if not ngp.has_device():
        raise RuntimeError("bench.py needs a CUDA device: ngp-b200 has no CPU fallback")
    rank, world, local = dist_setup()
    peaks = load_peaks()
    dev = torch.device("cuda", local)
    batch = FOX["batch"] * world   # weak scaling: 2^18 compacted samples per GPU and step
    K, W = args.steps, max(args.warmup, 3)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    tb = ngp.Testbed(device=dev)
    tb.use_current_stream()
    large = args.workload == "large"   # configs[4]: T = 2^22 (87.6 M parameters, 175 MB fp16: the table spills nothing but the collective is 7x larger)
    use_fox = os.path.isdir(FOX_DIR) and not os.environ.get("NGPB_BENCH_SYNTHETIC") and not large
    host_images = []
    if use_fox:
        # the reference's own scene (configs[2]): 50 JPEGs 1080x1920, OpenCV lens, aabb_scale 4; decoded on the host once,
        # kept in pinned memory for the end-to-end leg
        loader = importlib.import_module("instant-ngp-rendering_b200.loader")
        ds = loader.describe_dataset(FOX_DIR)
        tb.create_empty_nerf_dataset(len(ds["frames"]), aabb_scale=ds["aabb_scale"], is_hdr=ds["is_hdr"])
        for i, fr in enumerate(ds["frames"]):
            px = torch.from_numpy(np.ascontiguousarray(loader.load_frame_pixels(fr, ds))).pin_memory()
            h_, w_ = px.shape[:2]
            fx, fy = loader.finish_frame(fr, w_, h_)
            tb.nerf.training.set_image(i, px)
            cx, cy = fr["principal"]
            k1, k2, p1, p2 = fr["lens_params"][:4] if fr["lens_mode"] == 1 else (0.0, 0.0, 0.0, 0.0)
            tb.nerf.training.set_camera_intrinsics(i, fx=fx, fy=fy, cx=cx * w_, cy=cy * h_, k1=k1, k2=k2, p1=p1, p2=p2)
            tb.nerf.training.set_camera_extrinsics(i, fr["xform"], convert_to_ngp=False)
            host_images.append(px)
        tb.dataset, tb.up_dir, tb.mode = ds, ds["up"], tbm.MODE_NERF
        tb.nerf.training.near_distance = 0.1   # scripts/run.py default for real scenes
        data_name = "fox"
        dataset_desc = "data/nerf/fox of the reference (staged under data_stage/): %d JPEG views %dx%d, OpenCV lens, aabb_scale %d" % (len(host_images), w_, h_, ds["aabb_scale"])
    else:
        scene, poses, focal = syn.make_dataset(tb, n_images=FOX["n_images"], width=FOX["width"], height=FOX["height"], aabb_scale=FOX["aabb_scale"], amp_scale=FOX["amp_scale"])
        data_name = "synthetic"
        dataset_desc = f"{FOX['n_images']} views {FOX['width']}x{FOX['height']} RGBA8, aabb_scale {FOX['aabb_scale']}, analytic blob scene rendered on the GPU (data_stage/fox absent)"
    net_config = json.loads(json.dumps(syn.NERF_BASE_CONFIG))
    if large:
        net_config["encoding"]["log2_hashmap_size"] = 22
    tb.reload_network_from_json(net_config)
    if world > 1:
        tb.init_data_parallel()   # NCCL communicators inside the library; train() then runs the sharded step (csrc/dist.cu)
    step = lambda: tb.train(batch)

    # untimed: get past the dense-occupancy phase (first 256 steps) into steady state, then warm up
    for _ in range(FOX["pretrain_steps"]):
        step()
    for _ in range(W):
        step()
    sampler = ClockSampler(local)
    barrier(world)
    if rank == 0:
        sampler.start()
    e0, e1 = ev(), ev()
    compacted = marched = rays = 0
    e0.record()
    for _ in range(K):
        step()
        compacted += tb.measured_batch_size
        marched += tb.measured_batch_size_before_compaction
        rays += tb.rays_per_batch
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    clocks = sampler.stop() if rank == 0 else None
    ms = reduce_over_ranks(e0.elapsed_time(e1), world)
    value = compacted / (ms * 1e-3)     # counters are global (all-reduced) in a sharded run
    loss = tb.loss

    # phase breakdown of the same step (library phase timer = CUDA events between kernel groups)
    P = min(K, 32)
    tbm.profile_enable(True)
    pc = pm = 0
    for _ in range(P):
        step()
        pc += tb.measured_batch_size
        pm += tb.measured_batch_size_before_compaction
    torch.cuda.synchronize()
    prof = tbm.profile_collect()
    tbm.profile_enable(False)
    phases = nerf_phase_costs(prof, P, pc / P / world, pm / P / world, tb.n_params, peaks)
    timed = {k: v for k, v in phases.items() if "frac" in v}
    dom = max(timed, key=lambda k: timed[k]["ms_per_step"]) if timed else None

    if large:
        if rank == 0:
            dp = {k: v["ms_per_step"] for k, v in phases.items() if k.startswith("dp.") or k.startswith("optimizer")}
            emit_line({"metric": "large-scene NeRF train samples/s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
                       "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
                       "config": {"workload": "nerf_train_large_T2^22_batch2^18_per_gpu", "dataset": dataset_desc, "n_params": tb.n_params, "global_batch": batch,
                                  "parallelism": "1 GPU" if world == 1 else f"dp{world}: rays sharded; reduce-scatter of {tb.n_params} fp16 gradients, Adam on 1/{world} of them, all-gather of the fp16 weights"},
                       "samples_per_step": compacted / K, "loss": loss, "phases": phases, "dp": dp, "roofline": None, "gpu_launches": count_launches(prof, P, K), "clocks": clocks,
                       "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 20, "note": "device-timed; the e2e leg belongs to the default workload"}})
        return

    # end to end: the dataset comes from pinned host memory (re-uploaded inside the timed region through
    # set_nerf_image, as load_training_data does once per run), then K Testbed.train calls, each
    # reading its counters / loss back to the host.
    if not use_fox:
        for i in range(FOX["n_images"]):
            img = syn.linear_premul_to_srgb8(scene.render(poses[i], FOX["width"], FOX["height"], focal)).cpu().pin_memory()
            host_images.append(img)
    torch.cuda.synchronize()
    barrier(world)
    t0 = time.perf_counter()
    for i, img in enumerate(host_images):
        tb.set_nerf_image(i, img)
    e2e_samples = 0
    for _ in range(K):
        step()
        e2e_samples += tb.measured_batch_size
    _ = tb.loss
    torch.cuda.synchronize()
    e2e_s = reduce_over_ranks(time.perf_counter() - t0, world)
    h2d = sum(int(im.numel()) for im in host_images) / K
    del host_images

    # configs[3]: novel-view render of the trained model, 1920x1080, tiles sharded over the ranks
    rw, rh = 1920, 1080
    shard = (rank, world)
    if use_fox:
        # a pose between training views 0 and 1 (not a training pose), the views' own field of view on the landscape frame
        tb.set_camera_to_training_view(0)
        cam_a = np.array(tb._camera, dtype=np.float32)
        tb.set_camera_to_training_view(1)
        view = cam_a.copy()
        view[:, 3] = 0.5 * (cam_a[:, 3] + np.array(tb._camera, dtype=np.float32)[:, 3])
        focal = float(np.broadcast_to(tb._relative_focal_length, (2,))[1]) * rh
    else:
        view = syn.look_at_ngp(np.array([0.5, 0.5, 0.5]) + 1.4 * np.array([0.62, -0.55, 0.56]))  # not a training pose
    for _ in range(3):
        tb.render_nerf(rw, rh, view, focal, shard=shard)
    frames = 5
    barrier(world)
    r0, r1 = ev(), ev()
    r0.record()
    for _ in range(frames):
        rgba, depth = tb.render_nerf(rw, rh, view, focal, shard=shard)
    r1.record()
    torch.cuda.synchronize()
    render_ms = reduce_over_ranks(r0.elapsed_time(r1), world) / frames
    queries = reduce_over_ranks(float(tb.last_render_samples), world, op="sum")
    # end to end: this rank's tiles (256-pixel runs, interleaved over the ranks) to pinned host memory
    n_tiles_frame = rw * rh // 256
    host_tiles = torch.empty((len(range(rank, n_tiles_frame, world)), 256, 4), dtype=torch.float32).pin_memory()
    t0 = time.perf_counter()
    for _ in range(frames):
        rgba, depth = tb.render_nerf(rw, rh, view, focal, shard=shard)
        host_tiles.copy_(rgba.view(n_tiles_frame, 256, 4)[rank::world], non_blocking=True)
        torch.cuda.synchronize()
    render_e2e_ms = reduce_over_ranks((time.perf_counter() - t0) * 1e3, world) / frames
    coverage = float((rgba[..., 3] > 0.5).float().mean()) * world
    render = {"metric": "render Mrays/s", "value": rw * rh / (render_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_frame": render_ms, "resolution": [rw, rh],
              "queries_per_ray": queries / (rw * rh), "gqueries_per_s": queries / (render_ms * 1e-3) / 1e9, "coverage": coverage,
              "e2e": {"value": rw * rh / (render_e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "d2h_bytes_per_frame": rw * rh * 16 // world},
              "sharding": f"256-pixel tiles interleaved over {world} rank(s), no communication"}
    # per query 588 B of encoder traffic (L2-resident table: HBM is the floor) and 20 480 FLOP of MLP
    r_t = queries * NERF_MLP_FLOP_INFER / (render_ms * 1e-3) / 1e12 / world
    r_b = queries * NERF_GRID_BYTES_INFER / (render_ms * 1e-3) / 1e9 / world
    pk_t = peaks["bf16_tflops_sustained"] or peaks["bf16_tflops"]
    render["roofline"] = {"bound": "tensor", "achieved": r_t, "peak": pk_t, "unit": "TFLOP/s", "frac": r_t / pk_t, "traffic": None,
                          "gather": {"achieved": r_b, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": r_b / peaks["hbm_gbs"]},
                          "note": "whole frame (march + encoder + MLP + composite) against the MLP's tensor peak and the encoder's algorithmic bytes against the HBM copy peak, per GPU"}

    grid = grid_measure(ngp, min(K, 20), 3, rank, world, local)
    image = image_measure(ngp, min(K, 100), 3, rank, world, local)
    if rank != 0:
        return
    cpu = cpu_nerf_baseline(n_rays=256, steps=4, budget_s=20.0)
    gb = gpu_baseline() if (world == 1 and not os.environ.get("NGPB_BENCH_NO_BASELINE")) else None   # the switch exists for short ncu captures
    ref_train = (gb or {}).get("train", {}).get("samples_per_s")
    roof = None
    if dom:
        d = timed[dom]
        roof = {"bound": d["bound"], "kernel": dom, "achieved": d["achieved"], "peak": d["peak"], "unit": d["unit"], "frac": d["frac"],
                "traffic": load_traffic(dom), "peak_source": peaks["source"],
                "note": "largest roofline-bound kernel group of the training step by the phase timer (the ray march is issue-latency bound and overlapped, see phases.generate_training_samples); per-group figures in `phases`"}
    line = {
        "metric": "fox NeRF train samples/s", "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
        # the reference's own CUDA build on this GPU (measured in this run when baseline/_ref and the staged fox
        # are present, else the figure recorded in BASELINE.md from profiles/r02_reference_pyngp.json: 105.4 M samples/s)
        "vs_baseline": (value / ref_train) if (ref_train and use_fox) else ((value / 105.4e6) if use_fox else None),
        "dtype": "f16", "data": data_name,
        "config": {"workload": "nerf_train_fox_base.json_batch2^18" if use_fox else "nerf_train_fox_shape_base.json_batch2^18",
                   "dataset": dataset_desc,
                   "network": "configs/nerf/base.json (HashGrid L8 F4 T2^19 + 64-wide density/rgb MLPs), Ema(ExponentialDecay(Adam))",
                   "global_batch": batch, "pretrain_steps": FOX["pretrain_steps"],
                   "l2": "not flushed: a step streams the 283 MB optimizer state and gathers from 415 MB of images, both larger than the 126 MB L2",
                   "parallelism": "1 GPU" if world == 1 else f"dp{world}: rays sharded, NCCL all-reduce of {tb.n_params} fp16 gradients per step"},
        "samples_per_step": compacted / K, "marched_samples_per_step": marched / K, "rays_per_step": rays / K, "loss": loss,
        "phases": phases,
        "overlap": {"sum_of_phases_ms": sum(v["ms_per_step"] for v in phases.values()), "ms_per_step": ms / K,
                    "note": "phases are timed serially (prefetch off under the phase timer); the step itself overlaps the next step's ray march with backward + optimizer"},
        "roofline": roof,
        "cpu_baseline": {"value": cpu["samples_per_s"], "unit": "samples/s", "cores": cpu["cores"], "kind": "port", "sample": cpu["sample"]},
        "e2e": {"value": e2e_samples / e2e_s, "unit": "samples/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 20,
                "note": "dataset uploaded from pinned host memory inside the timed region, then K Testbed.train calls with their counter read-backs"},
        "gpu_launches": count_launches(prof, P, K),
        "gpu_baseline": gb,
        "render": render, "grid": grid, "image": image, "clocks": clocks,
    }
    emit_line(line)


def run_grid(args):
    ngp = importlib.import_module("instant-ngp-rendering_b200")
    if not ngp.has_device():
        raise RuntimeError("bench.py needs a CUDA device: ngp-b200 has no CPU fallback")
    rank, world, local = dist_setup()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    g = grid_measure(ngp, args.steps, args.warmup, rank, world, local)
    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        return
    cpu_v, cpu_dt, cores = cpu_grid_baseline(1 << 18)
    line = {"metric": g["metric"], "value": g["value"], "unit": g["unit"], "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": g["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
            "config": g["config"], "kernels_ms": g["kernels_ms"], "roofline": g["roofline"],
            "cpu_baseline": {"value": cpu_v, "unit": "Gpts/s", "cores": cores, "kind": "port", "sample": "2^18 uniform points fwd+bwd (oracle, OpenMP)"},
            "e2e": g["e2e"], "gpu_launches": g["gpu_launches"], "clocks": clocks}
    emit_line(line)


def run_reference_arm(args):
    """--impl reference. The reference has no CPU implementation of this path and its CUDA build
    needs cmake + a GPU (DESIGN.md "Reference arm"), so the arm times the oracle port of the same
    workload on all host threads, on a bounded sample per step. Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    if args.workload == "grid":
        sample = 1 << 18
        for _ in range(min(args.warmup, 2)):
            cpu_grid_baseline(1 << 14)
        t_total, cores = 0.0, 1
        for _ in range(args.steps):
            _, dt, cores = cpu_grid_baseline(sample)
            t_total += dt
            if t_total > 60:
                break
        value = sample * args.steps / t_total / 1e9
        line = {"impl": "reference", "metric": "hash-enc fwd+bwd Gpts/s", "value": value, "unit": "Gpts/s", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps * (GRID_N / sample), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f16", "data": "synthetic", "config": {"workload": "grid_L16F2_T19_2^20pts", "sample": "2^18 of 2^20 points per step"},
                "cpu_baseline": {"value": value, "unit": "Gpts/s", "cores": cores, "kind": "port", "sample": "2^18 uniform points, fwd+bwd, OpenMP"},
                "e2e": {"value": value, "unit": "Gpts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        emit_line(line)
        return
    cpu = cpu_nerf_baseline(n_rays=256, steps=max(args.steps, 1), warmup=min(args.warmup, 1), budget_s=90.0)
    value = cpu["samples_per_s"]
    per_step = cpu["samples"] / max(cpu["steps_done"], 1)
    line = {"impl": "reference", "metric": "fox NeRF train samples/s", "value": value, "unit": "samples/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * cpu["seconds"] / max(cpu["steps_done"], 1) * (FOX["batch"] / max(per_step, 1)),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f16", "data": cpu["data"],
            "config": {"workload": "nerf_train_fox_base.json_batch2^18" if cpu["data"] == "fox" else "nerf_train_fox_shape_base.json_batch2^18", "sample": cpu["sample"],
                       "note": "ms_per_step is the measured time per oracle step scaled to the 2^18-sample batch"},
            "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cpu["cores"], "kind": "port", "sample": cpu["sample"]},
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit_line(line)


_REAL_STDOUT = None


def claim_stdout():
    """stdout carries ONE JSON line. Libraries print there too (NCCL's version banner under NCCL_DEBUG=VERSION, which
    the GPU box sets, is a plain printf): point fd 1 at stderr for the run and keep the real stdout for the line."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_REAL_STDOUT, data)


def main():
    claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", default="nerf", choices=["nerf", "grid", "large"])
    ap.add_argument("--impl", default="ours")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    if args.workload == "grid":
        run_grid(args)
    else:
        run_nerf(args)
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
