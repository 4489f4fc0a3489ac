#!/usr/bin/env bash
# Snapshots written / read by the library's host C++ against the reference's own build, on fox (1000 steps each side):
# reference reads the C++-written file, the C++ reader reads the reference-written file; PSNR protocol on both.
mkdir -p gpurun_out
timeout 200 python baseline/run_pyngp.py --impl ours --native 1 --what fox --steps 1000 --eval 1 --frames 2 --save /tmp/ours_native.ingp --out gpurun_out/r2_native_ours.json > gpurun_out/r2_native_ours.log 2>&1; echo "ours native rc=$?"
timeout 200 python baseline/run_pyngp.py --impl reference --what fox --steps 1000 --eval 1 --frames 2 --save /tmp/ref1000.ingp --out gpurun_out/r2_native_ref.json > gpurun_out/r2_native_ref.log 2>&1; echo "ref rc=$?"
timeout 200 python baseline/run_pyngp.py --impl reference --what fox --load /tmp/ours_native.ingp --frames 2 --out gpurun_out/r2_ref_loads_native.json > gpurun_out/r2_ref_loads_native.log 2>&1; echo "ref<-native rc=$?"
timeout 200 python baseline/run_pyngp.py --impl ours --native 1 --what fox --load /tmp/ref1000.ingp --frames 2 --out gpurun_out/r2_native_loads_ref.json > gpurun_out/r2_native_loads_ref.log 2>&1; echo "native<-ref rc=$?"
python - <<PY
import json
for f in ("r2_native_ours", "r2_native_ref", "r2_ref_loads_native", "r2_native_loads_ref"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f))["fox"]
        if "eval" in d: print(f, "step", d.get("training_step"), "psnr", round(d["eval"]["psnr_mean"], 3))
        else: print(f, [(c["step"], round(c["psnr_mean"], 3)) for c in d["checkpoints"]], "load_s", round(d["load_s"], 2))
    except Exception as e:
        print(f, "failed", e)
PY
