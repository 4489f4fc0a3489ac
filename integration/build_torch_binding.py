"""Builds the reference's own PyTorch extension on top of this library.

tiny-cuda-nn's torch binding (tcnn/bindings/torch/tinycudann/bindings.cpp) talks to the library through tcnn::cpp::Module only.
This script compiles that file UNMODIFIED, from where it lies under /root/reference, together with
integration/tcnn_cpp_api_b200.cpp, and links libngp_b200.so:

    integration/_build/torch/tinycudann_bindings/_90_C.<abi>.so     (the slot name tinycudann/modules.py looks for: the
                                                                     highest compute capability in its list not above the GPU's)
    baseline/_ref/tinycudann_pkg/tinycudann/{__init__,modules}.py    (an install of the reference's Python package - git-ignored
                                                                     like the rest of baseline/_ref, it travels to the GPU box)

With both directories on sys.path, `import tinycudann as tcnn; tcnn.NetworkWithInputEncoding(...)` is the reference's Python
and binding code running on the sm_100a kernels. Only where /root/reference exists; g++ only (no CUDA kernels are compiled).
"""
import os
import shutil
import subprocess
import sys
import sysconfig

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(os.environ.get("NGP_REFERENCE", "/root/reference"), "dependencies", "tiny-cuda-nn")
OUT_DIR = os.path.join(ROOT, "integration", "_build", "torch", "tinycudann_bindings")
PKG_DIR = os.path.join(ROOT, "baseline", "_ref", "tinycudann_pkg", "tinycudann")


def main():
    src = os.path.join(REF, "bindings", "torch", "tinycudann", "bindings.cpp")
    if not os.path.exists(src):
        print("build_torch_binding: reference absent, keeping prebuilt files", file=sys.stderr)
        return 0
    import torch
    from torch.utils import cpp_extension as ce
    shim = os.path.join(ROOT, "integration", "tcnn_cpp_api_b200.cpp")
    lib = os.path.join(ROOT, "instant-ngp-rendering_b200", "libngp_b200.so")
    out = os.path.join(OUT_DIR, "_90_C" + sysconfig.get_config_var("EXT_SUFFIX"))
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(PKG_DIR, exist_ok=True)
    for name in ("__init__.py", "modules.py"):
        shutil.copyfile(os.path.join(REF, "bindings", "torch", "tinycudann", name), os.path.join(PKG_DIR, name))
    deps = [src, shim, os.path.join(ROOT, "include", "ngp_b200.h")]
    if os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return 0
    try:
        inc = ce.include_paths("cuda")
    except TypeError:
        inc = ce.include_paths(cuda=True)
    cmd = ["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-w", "-DTORCH_EXTENSION_NAME=_90_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
           "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)]
    cmd += ["-I" + i for i in inc] + ["-I" + sysconfig.get_paths()["include"], "-I" + os.path.join(REF, "include"), "-I" + os.path.join(REF, "dependencies"),
                                     "-I" + os.path.join(ROOT, "include"), "-I/usr/local/cuda/include"]
    cmd += [src, shim, "-L" + os.path.dirname(lib), "-lngp_b200", "-Wl,-rpath,$ORIGIN/../../../../instant-ngp-rendering_b200",
            "-L" + os.path.join(os.path.dirname(torch.__file__), "lib"), "-ltorch", "-ltorch_cpu", "-ltorch_cuda", "-lc10", "-lc10_cuda", "-ltorch_python",
            "-L/usr/local/cuda/lib64", "-lcudart", "-o", out]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stderr[-4000:])
        return 1
    print("build_torch_binding:", out, file=sys.stderr)
    return 0


if __name__ == "__main__":
    sys.exit(main())
