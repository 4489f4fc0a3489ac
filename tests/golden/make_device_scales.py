"""Writes tests/golden/device_level_scales.json: the per-level grid scales as the GPU computes them (exp2f on the device,
as tcnn's kernel_grid does, common_device.h:709-718) for the grids of tests/ref_gpu_cases.py. The host's exp2f differs from
the device's in the last bit on some levels, and a CPU test that compares the oracle with the reference's GPU kernel outputs
(tests/test_ref_gpu_fixture_cpu.py) must interpolate on the scales those kernels used. Run on a GPU box:
    python tests/golden/make_device_scales.py"""
import importlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_gpu_cases as rc

ngp = importlib.import_module("instant-ngp-rendering_b200")
out = {}
cases = {name: (d, enc) for name, (d, enc, _) in rc.GRID_CASES.items()}
cases["nerfnet"] = (3, rc.nerfnet_case()[0]["encoding"])
for name, (n_dims, enc) in cases.items():
    _, scales, _ = ngp.Encoding(n_dims, enc).level_table()
    out[name] = [float(s) for s in scales]
dst = os.path.join(ROOT, "gpurun_out", "device_level_scales.json")
os.makedirs(os.path.dirname(dst), exist_ok=True)
json.dump(out, open(dst, "w"), indent=1)
print(out)
