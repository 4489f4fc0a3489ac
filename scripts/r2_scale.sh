#!/usr/bin/env bash
# round-2 scaling runs on N GPUs of one box: data-parallel check, the default bench, the large (T = 2^22) workload
N=${1:-8}
P=$((29600 + N))
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $P scripts/dp_check.py 2>&1 | grep -E "^\{" | tee gpurun_out/r2_dp_check_n$N.json
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((P+20)) bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "bench rc=$?"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((P+40)) bench.py --gpus $N --steps 20 --warmup 5 --workload large > gpurun_out/r2_large_n$N.json 2> gpurun_out/r2_large_n$N.err; echo "large rc=$?"
python - <<PY
import json
for f in ("gpurun_out/r2_bench_n$N.json", "gpurun_out/r2_large_n$N.json"):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f, {k: d[k] for k in ["value", "ms_per_step", "n_gpus"]})
        if "render" in d: print("  render", d["render"]["ms_per_frame"], d["render"]["value"], d["render"]["queries_per_ray"], "e2e", d["render"]["e2e"]["value"])
        print("  ", {k: round(v.get("ms_per_step"), 4) for k, v in d["phases"].items()})
    except Exception as e:
        print(f, "failed", e)
PY
