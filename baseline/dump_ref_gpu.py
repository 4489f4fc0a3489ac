"""Runs the reference's own GPU kernels (baseline/_ref/ref_harness) on the seeded inputs of tests/ref_gpu_cases.py
and writes their outputs as the committed golden vectors tests/golden/ref_gpu_vectors.npz, plus the reference's
kernel timings on the benchmark shapes (gpurun_out/ref_gpu_timings.json). Run on the GPU box:

    gpurun -- python baseline/dump_ref_gpu.py          # -> gpurun_out/ref_gpu_vectors.npz, ref_gpu_timings.json

then copy gpurun_out/ref_gpu_vectors.npz to tests/golden/. Dense gradients are stored sparsely (index + value of
the non-zero entries) to keep the fixture small.
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_gpu_cases as rc  # noqa: E402

HARNESS = os.path.join(ROOT, "baseline", "_ref", "ref_harness")
OUT = os.path.join(ROOT, "gpurun_out")


def run(mode, cfg, files, outputs):
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "cfg.json"), "w") as f:
            json.dump(cfg, f)
        for name, arr in files.items():
            np.ascontiguousarray(arr).tofile(os.path.join(d, name))
        res = subprocess.run([HARNESS, mode, d], capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("ref_harness %s failed: %s%s" % (mode, res.stdout, res.stderr))
        got = {}
        for name, dtype in outputs.items():
            p = os.path.join(d, name)
            if os.path.exists(p):
                got[name] = np.fromfile(p, dtype=dtype)
        t = os.path.join(d, "timing.json")
        timing = json.load(open(t)) if os.path.exists(t) else None
        return got, timing


def sparse(dense):
    idx = np.nonzero(dense.view(np.uint16) & 0x7FFF)[0].astype(np.uint32)
    return idx, dense[idx]


def dump_vectors():
    vec = {}
    for name in rc.GRID_CASES:
        n_dims, enc, x, params, dy = rc.grid_case(name)
        got, _ = run("grid", {"n_dims": n_dims, "encoding": enc}, {"x.bin": x, "params.bin": params, "dy.bin": dy}, {"y.bin": np.float16, "dparams.bin": np.float16})
        vec["grid/%s/y" % name] = got["y.bin"].reshape(len(x), -1)
        vec["grid/%s/dparams_idx" % name], vec["grid/%s/dparams_val" % name] = sparse(got["dparams.bin"])
        print("grid", name, "nonzero gradients", len(vec["grid/%s/dparams_idx" % name]), flush=True)
    for name in rc.MLP_CASES:
        n_in, n_out, cfg, x, w, dy = rc.mlp_case(name)
        got, _ = run("mlp", {"n_in": n_in, "n_out": n_out, "network": cfg}, {"x.bin": x, "params.bin": w, "dy.bin": dy},
                     {"y.bin": np.float16, "dx.bin": np.float32, "dparams.bin": np.float16})
        vec["mlp/%s/y" % name] = got["y.bin"].reshape(len(x), 16)
        vec["mlp/%s/dx" % name] = got["dx.bin"].reshape(len(x), n_in)
        vec["mlp/%s/dparams" % name] = got["dparams.bin"]
    cfg, x, params, dy = rc.nerfnet_case()
    got, _ = run("nerfnet", cfg, {"x.bin": x, "params.bin": params, "dy.bin": dy},
                 {"y.bin": np.float16, "y_inference.bin": np.float16, "density.bin": np.float16, "dparams.bin": np.float16})
    n_mlp = rc.mlp_n_params(32, 1) + rc.mlp_n_params(32, 2)
    vec["nerfnet/y"] = got["y.bin"].reshape(len(x), 16)[:, :4]
    vec["nerfnet/y_inference"] = got["y_inference.bin"].reshape(len(x), 16)[:, :4]
    vec["nerfnet/density"] = got["density.bin"].reshape(len(x), 16)[:, :1]
    vec["nerfnet/dparams_mlp"] = got["dparams.bin"][:n_mlp]
    vec["nerfnet/dparams_grid_idx"], vec["nerfnet/dparams_grid_val"] = sparse(got["dparams.bin"][n_mlp:])
    cfg, w, grads = rc.adam_case()
    got, info = run("adam", cfg, {"w.bin": w, "g.bin": grads}, {"w32.bin": np.float32, "w16.bin": np.float16, "ema.bin": np.float16})
    vec["adam/w32"], vec["adam/w16"], vec["adam/ema"] = got["w32.bin"], got["w16.bin"], got["ema.bin"]
    vec["adam/learning_rate"] = np.array([info["learning_rate"]], dtype=np.float32)
    np.savez_compressed(os.path.join(OUT, "ref_gpu_vectors.npz"), **vec)
    print("wrote", os.path.join(OUT, "ref_gpu_vectors.npz"), os.path.getsize(os.path.join(OUT, "ref_gpu_vectors.npz")), "bytes", flush=True)



def time_kernels(iters=20, out_path=None):
    """The reference's kernels on the benchmark shapes; returns the timings dict."""
    timings = {}
    g = np.random.default_rng(1337)
    for name, n in (("L16F2", 1 << 20), ("L8F4", 1 << 20)):
        n_dims, enc, _ = rc.GRID_CASES[name]
        x = g.random((n, n_dims), dtype=np.float32)
        params = (g.standard_normal(rc.grid_n_params(n_dims, enc)) * 0.1).astype(np.float16)
        dy = g.standard_normal((n, 32)).astype(np.float16)
        _, t = run("grid", {"n_dims": n_dims, "encoding": enc, "time_iters": iters}, {"x.bin": x, "params.bin": params, "dy.bin": dy}, {})
        f, b = float(np.median(t["fwd_ms"])), float(np.median(t["bwd_ms"]))
        timings["grid_" + name] = {"n": n, "fwd_ms": f, "bwd_ms": b, "inference_ms": float(np.median(t["inference_ms"])), "gpts_per_s_fwd_bwd": n / (f + b) / 1e6}
        print("grid", name, timings["grid_" + name], flush=True)
    n = 1 << 18
    cfg = {"encoding": rc.GRID_L8F4, "dir_encoding": rc.NERF_DIR, "network": rc.NERF_DENSITY, "rgb_network": rc.NERF_RGB, "time_iters": iters}
    x = g.random((n, 7), dtype=np.float32)
    params = np.concatenate([rc.mlp_weights(g, 32, 1), rc.mlp_weights(g, 32, 2), (g.standard_normal(rc.grid_n_params(3, rc.GRID_L8F4)) * 0.1).astype(np.float16)])
    dy = np.zeros((n, 16), dtype=np.float16)
    dy[:, :4] = (g.standard_normal((n, 4)) * 0.1).astype(np.float16)
    _, t = run("nerfnet", cfg, {"x.bin": x, "params.bin": params, "dy.bin": dy}, {})
    timings["nerfnet_2e18"] = {"n": n, "fwd_bwd_ms": float(np.median(t["fwd_bwd_ms"])), "inference_ms": float(np.median(t["inference_ms"]))}
    print("nerfnet", timings["nerfnet_2e18"], flush=True)
    n_in, n_out, mcfg, _, w, _ = rc.mlp_case("rgb_32_2")
    xm = g.uniform(-1, 1, size=(n, n_in)).astype(np.float32)
    dym = np.zeros((n, 16), dtype=np.float16)
    dym[:, :3] = (g.standard_normal((n, 3)) * 0.1).astype(np.float16)
    _, t = run("mlp", {"n_in": n_in, "n_out": n_out, "network": mcfg, "time_iters": iters}, {"x.bin": xm, "params.bin": w, "dy.bin": dym}, {})
    timings["mlp_32_64x2_3_2e18"] = {"n": n, "fwd_ms": float(np.median(t["fwd_ms"])), "bwd_ms": float(np.median(t["bwd_ms"])), "inference_ms": float(np.median(t["inference_ms"]))}
    print("mlp", timings["mlp_32_64x2_3_2e18"], flush=True)
    acfg = dict(rc.adam_case()[0], n=12865536, steps=1, time_iters=iters, layers=[[64, 32], [16, 64], [64, 32], [64, 64], [16, 64]])
    _, t = run("adam", acfg, {"w.bin": np.zeros(acfg["n"], np.float32), "g.bin": (g.standard_normal(acfg["n"])).astype(np.float16)}, {})
    timings["adam_12.9M"] = {"n": acfg["n"], "step_ms": float(np.median(t["step_ms"]))}
    print("adam", timings["adam_12.9M"], flush=True)
    if out_path:
        with open(out_path, "w") as f:
            json.dump(timings, f, indent=1)
    return timings


def main():
    os.makedirs(OUT, exist_ok=True)
    if "--timings-only" not in sys.argv:
        dump_vectors()
    time_kernels(out_path=os.path.join(OUT, "ref_gpu_timings.json"))


if __name__ == "__main__":
    main()
