"""Builds libngp_b200.so (CUDA kernels + C ABI, sm_100a only) and the CPU oracle library.

Run as `python instant-ngp-rendering_b200/build.py` or through `__graft_entry__.build()`.
nvcc cross-compiles without a GPU; objects are cached under build/ by source mtime.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
BUILD = os.path.join(ROOT, "build")
LIB_PATH = os.path.join(PKG_DIR, "libngp_b200.so")
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_LIB = os.path.join(ORACLE_DIR, "liboracle.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr", "--extended-lambda",
    "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function,-ffp-contract=off",
    "-Xptxas", "-v" if os.environ.get("NGPB_PTXAS_V") else "-O3",
]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
        raise RuntimeError("build step failed: " + " ".join(cmd[:3]))
    if os.environ.get("NGPB_PTXAS_V"):
        sys.stderr.write(res.stderr)
    return res


def build_cuda(verbose=False):
    os.makedirs(BUILD, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(ROOT, "include", "ngp_b200.h"))
    sources = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    objs, jobs = [], []
    for s in sources:
        src = os.path.join(CSRC, s)
        obj = os.path.join(BUILD, s[:-3] + ".o")
        objs.append(obj)
        if _newer(obj, [src] + headers):
            # NeRF ray set-up / stepping must give the same bits as the host oracle: no FMA contraction
            extra = ["-fmad=false"] if s.startswith("nerf_") else []
            jobs.append([NVCC] + NVCC_FLAGS + extra + ["-c", src, "-o", obj])
    if jobs:
        if verbose:
            print(f"[build] compiling {len(jobs)} CUDA translation unit(s) for sm_100a")
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(_run, jobs))
    if jobs or _newer(LIB_PATH, objs):
        _run([NVCC, "-shared", "-o", LIB_PATH] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart_static", "-lnvjpeg_static", "-lculibos", "-lz", "-ldl", "-lpthread", "-lrt"])
        if verbose:
            print(f"[build] linked {LIB_PATH}")
    return LIB_PATH


def build_oracle(verbose=False):
    srcs = sorted(os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".cpp", ".c")))
    hdrs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith(".h")]
    if not srcs:
        return None
    if _newer(ORACLE_LIB, srcs + hdrs):
        _run(["g++", "-O3", "-march=native", "-std=c++17", "-fopenmp", "-ffp-contract=off", "-fPIC", "-shared",
              "-o", ORACLE_LIB] + srcs)
        if verbose:
            print(f"[build] built {ORACLE_LIB}")
    return ORACLE_LIB


def build_ref(verbose=False):
    """oracle/_ref/libref_host.so: the reference's own header functions compiled for the host
    (oracle/ref_host.cu, oracle/build_ref.sh). Only where /root/reference exists; a checker, like the oracle."""
    ref = os.environ.get("NGP_REFERENCE", "/root/reference")
    out = os.path.join(ORACLE_DIR, "_ref", "libref_host.so")
    src = os.path.join(ORACLE_DIR, "ref_host.cu")
    if not os.path.isdir(os.path.join(ref, "include", "neural-graphics-primitives")):
        return out if os.path.exists(out) else None
    if _newer(out, [src, os.path.join(ORACLE_DIR, "build_ref.sh")]):
        _run(["bash", os.path.join(ORACLE_DIR, "build_ref.sh")])
        if verbose:
            print(f"[build] built {out}")
    return out


def build_baseline(verbose=False):
    """baseline/_ref: the UNMODIFIED reference (instant-ngp + tiny-cuda-nn) compiled headless for sm_100 by
    baseline/build_ref.sh (direct nvcc recipe; ~25 min of CPU from scratch), the GPU baseline and GPU oracle of
    DESIGN.md section 4. Relinked here when its libraries exist; a from-scratch build needs NGPB_BUILD_REFERENCE=1
    (or run baseline/build_ref.sh). Never part of the product."""
    base = os.path.join(ROOT, "baseline")
    ref = os.environ.get("NGP_REFERENCE", "/root/reference")
    if not os.path.isdir(os.path.join(ref, "src")):
        return
    if os.path.exists(os.path.join(base, "_ref", "obj", "libtcnn_ref.a")) and os.path.exists(os.path.join(base, "_ref", "obj", "libngp_ref.a")):
        _run(["bash", os.path.join(base, "build_harness.sh")])
        if verbose:
            print("[build] baseline/_ref up to date (ref_harness, pyngp)")
    elif os.environ.get("NGPB_BUILD_REFERENCE"):
        _run(["bash", os.path.join(base, "build_ref.sh")])
    elif verbose:
        print("[build] baseline/_ref not built (NGPB_BUILD_REFERENCE=1 or baseline/build_ref.sh builds the reference: ~25 min)")


def stage_data(verbose=False):
    """Copies the reference's sample inputs that bench.py measures on (data/nerf/fox, data/image/albert.exr) to the
    git-ignored data_stage/, so that they travel to the GPU box; a no-op where /root/reference does not exist."""
    import shutil
    ref_data = "/root/reference/data"
    stage = os.path.join(ROOT, "data_stage")
    if not os.path.isdir(ref_data):
        return
    fox_src, fox_dst = os.path.join(ref_data, "nerf", "fox"), os.path.join(stage, "fox")
    if os.path.isdir(fox_src) and not os.path.exists(os.path.join(fox_dst, "transforms.json")):
        shutil.copytree(fox_src, fox_dst, dirs_exist_ok=True)
        if verbose:
            print(f"[build] staged {fox_dst}")
    alb_src, alb_dst = os.path.join(ref_data, "image", "albert.exr"), os.path.join(stage, "albert.exr")
    if os.path.exists(alb_src) and not os.path.exists(alb_dst):
        os.makedirs(stage, exist_ok=True)
        shutil.copyfile(alb_src, alb_dst)
        if verbose:
            print(f"[build] staged {alb_dst}")


def build_examples(verbose=False):
    """examples/ngp_host: a C++ host of the C ABI with no Python in the process (load scene, load config file, train, snapshot, frame)."""
    src = os.path.join(ROOT, "examples", "ngp_host.cpp")
    exe = os.path.join(ROOT, "examples", "ngp_host")
    if not os.path.exists(src):
        return None
    if _newer(exe, [src, LIB_PATH, os.path.join(ROOT, "include", "ngp_b200.h")]):
        _run(["g++", "-std=c++17", "-O2", src, "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include", "-L" + PKG_DIR, "-lngp_b200",
              "-L/usr/local/cuda/lib64", "-lcudart", "-Wl,-rpath,$ORIGIN/../instant-ngp-rendering_b200", "-Wl,-rpath,/usr/local/cuda/lib64", "-o", exe])
        if verbose:
            print(f"[build] built {exe}")
    return exe


def build_integration(verbose=False):
    """integration/_build/tcnn_caller_{b200,ref}: one caller of the reference's tcnn::cpp facade linked against this library (through
    integration/tcnn_cpp_api_b200.cpp) and against the reference's own tcnn build. Needs the reference's headers: only where it exists."""
    script = os.path.join(ROOT, "integration", "build.sh")
    if os.path.exists(script):
        _run(["bash", script])
    torch_binding = os.path.join(ROOT, "integration", "build_torch_binding.py")
    if os.path.exists(torch_binding):
        _run([sys.executable, torch_binding])   # the reference's bindings.cpp, unmodified, over the same shim
        if verbose:
            print("[build] integration/_build is up to date")


def build_all(verbose=False):
    build_cuda(verbose)
    build_oracle(verbose)
    build_ref(verbose)
    build_baseline(verbose)
    build_examples(verbose)
    build_integration(verbose)
    stage_data(verbose)


if __name__ == "__main__":
    build_all(verbose=True)
