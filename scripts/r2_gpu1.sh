#!/usr/bin/env bash
# round 2, GPU call 1: parity suite after the deterministic-compositing change, the reference's kernels
# (golden dumps + timings), the reference and this repo through the same pyngp protocol, snapshot cross-loading
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r2_tests1.log 2>&1; echo "tests rc=$?"
timeout 900 python baseline/dump_ref_gpu.py > gpurun_out/r2_dump.log 2>&1; echo "dump rc=$?"
timeout 900 python baseline/run_pyngp.py --impl reference --what fox,albert --save /tmp/ref_fox.ingp --out gpurun_out/r2_ref_pyngp.json > gpurun_out/r2_ref_pyngp.log 2>&1; echo "ref rc=$?"
timeout 900 python baseline/run_pyngp.py --impl ours --what fox,albert --save /tmp/ours_fox.ingp --out gpurun_out/r2_ours_pyngp.json > gpurun_out/r2_ours_pyngp.log 2>&1; echo "ours rc=$?"
timeout 600 python baseline/run_pyngp.py --impl ours --what fox --load /tmp/ref_fox.ingp --out gpurun_out/r2_ours_loads_ref.json > gpurun_out/r2_ours_loads_ref.log 2>&1; echo "ours<-ref rc=$?"
timeout 600 python baseline/run_pyngp.py --impl reference --what fox --load /tmp/ours_fox.ingp --out gpurun_out/r2_ref_loads_ours.json > gpurun_out/r2_ref_loads_ours.log 2>&1; echo "ref<-ours rc=$?"
python baseline/strip_snapshot.py /tmp/ref_fox.ingp gpurun_out/ref_fox_step5000.snapshot.json
python baseline/strip_snapshot.py /tmp/ours_fox.ingp gpurun_out/ours_fox_step5000.snapshot.json
tail -3 gpurun_out/r2_tests1.log; tail -5 gpurun_out/r2_dump.log; tail -4 gpurun_out/r2_ref_pyngp.log; tail -4 gpurun_out/r2_ours_pyngp.log; tail -2 gpurun_out/r2_ours_loads_ref.log; tail -2 gpurun_out/r2_ref_loads_ours.log
