"""Prints the kernels of a traced training step (gpurun_out/timeline_train.json from scripts/timeline_train.py): stream, start
offset and duration in microseconds. Diagnosis only."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ev = json.load(open(os.path.join(ROOT, "gpurun_out", "timeline_train.json")))["traceEvents"]
k = sorted([e for e in ev if e.get("cat") in ("kernel", "gpu_memset", "gpu_memcpy")], key=lambda e: e["ts"])
t0 = k[0]["ts"]
# second step only: from the 2nd occurrence of the optimizer kernel to the 3rd
opt = [i for i, e in enumerate(k) if "adam_ema_step" in e["name"]]
lo, hi = (opt[0] + 1, opt[1] + 1) if len(opt) > 1 else (0, len(k))
base = k[lo]["ts"]
for e in k[lo:hi + 6]:
    print("%8.1f %7.1f  s%-3s %s" % (e["ts"] - base, e["dur"], e["args"].get("stream", "?"), e["name"][:70]))
cpu = sorted([e for e in ev if e.get("cat") == "cuda_runtime" and "Synchronize" in e.get("name", "")], key=lambda e: e["ts"])
print("host syncs in trace:", len(cpu), "mean dur us", sum(e["dur"] for e in cpu) / max(len(cpu), 1))
