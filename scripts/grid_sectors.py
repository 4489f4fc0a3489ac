"""profiles/grid_sectors.json from an ncu capture of `bench.py --workload grid`:
  ncu --metrics lts__t_sectors.sum,lts__t_sectors_op_read.sum,lts__t_sectors_op_red.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum \
      --clock-control none -k regex:grid_ --csv --log-file gpurun_out/grid_sectors.csv python bench.py --workload grid --steps 2 --warmup 1
Usage: python scripts/grid_sectors.py gpurun_out/grid_sectors.csv profiles/grid_sectors.json   (the LAST launch of every kernel is taken)"""
import csv
import json
import re
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]
out = {}
for r in rows[1:]:
    d = dict(zip(h, r))
    name = re.sub(r"^void |ngpb::|<.*|\(.*", "", d["Kernel Name"])
    v = float(d["Metric Value"].replace(",", ""))
    unit = d["Metric Unit"]
    v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1.0, "ms": 1e3, "ns": 1e-3}.get(unit, 1.0)
    out.setdefault(name, {"id": d["ID"]})
    if out[name]["id"] != d["ID"]:
        out[name] = {"id": d["ID"]}
    out[name][d["Metric Name"]] = v
res = {"points": 1 << 20, "_source": "ncu --clock-control none on `bench.py --workload grid` (config 1: L16 F2 T=2^19, 2^20 uniform points, L2 flushed before the step); last launch of each kernel"}
for name, m in out.items():
    if not name.startswith("grid_"):
        continue
    res[name] = {"lts_sectors": m.get("lts__t_sectors.sum"), "lts_sectors_read": m.get("lts__t_sectors_op_read.sum"), "lts_sectors_red": m.get("lts__t_sectors_op_red.sum"),
                 "dram_bytes": (m.get("dram__bytes_read.sum", 0) or 0) + (m.get("dram__bytes_write.sum", 0) or 0), "time_us_under_ncu": m.get("gpu__time_duration.sum")}
json.dump(res, open(sys.argv[2], "w"), indent=1)
print(json.dumps(res, indent=1))
