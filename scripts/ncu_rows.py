"""Per-kernel rows out of an `ncu --set full` report: ncu -i REP --page raw --csv -> JSON (profiles/rNN_*_ncu_rows.json).

Usage: python scripts/ncu_rows.py gpurun_out/r02_nerf_step.ncu-rep profiles/r02_nerf_ncu_rows.json
"""
import csv
import io
import json
import re
import subprocess
import sys

COLS = {
    "time_us": ("gpu__time_duration.sum", None),
    "warps_active_pct": ("sm__warps_active.avg.pct_of_peak_sustained_active", None),
    "tensor_pipe_pct": ("sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active", None),
    "sm_throughput_pct": ("sm__throughput.avg.pct_of_peak_sustained_elapsed", None),
    "lts_throughput_pct": ("lts__throughput.avg.pct_of_peak_sustained_elapsed", None),
    "dram_throughput_pct": ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "FBSP.TriageCompute.dram__throughput.avg.pct_of_peak_sustained_elapsed"),
    "dram_bytes_read": ("dram__bytes_read.sum", None),
    "dram_bytes_write": ("dram__bytes_write.sum", None),
    "grid": ("launch__grid_size", None),
    "block": ("launch__block_size", None),
    "regs": ("launch__registers_per_thread", None),
}
SCALE = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    result = []
    for r in body:
        name = re.sub(r"\(.*", "", r[col["Kernel Name"]])
        d = {"kernel": name}
        for key, (a, b) in COLS.items():
            c = col.get(a, col.get(b) if b else None)
            if c is None or r[c] in ("", "n/a"):
                d[key] = None
                continue
            v = float(r[c].replace(",", ""))
            d[key] = round(v * SCALE.get(units[c], 1.0), 3)
        if d["dram_bytes_read"] is not None and d["dram_bytes_write"] is not None:
            d["dram_traffic_bytes"] = d["dram_bytes_read"] + d["dram_bytes_write"]
        result.append(d)
    doc = {"source": rep, "how": "ncu --set full --clock-control none --import-source on (scripts/r2_ncu.sh): one steady-state training step on fox, 2^18 samples; "
           "times are per launch under the profiler (cold caches, serialised) - shares, not bench values", "kernels": result}
    json.dump(doc, open(out, "w"), indent=1)
    for d in result:
        print("%-60s %8.1f us  warps %5s  tensor %5s  lts %5s  dram %5s  traffic %s" % (d["kernel"][:60], d["time_us"], d["warps_active_pct"], d["tensor_pipe_pct"],
              d["lts_throughput_pct"], d["dram_throughput_pct"], d.get("dram_traffic_bytes")))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
