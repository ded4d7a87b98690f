"""tools/dram_traffic.py -- turns an ncu CSV (--metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum
over every launch of tools/prof_render.py <workload> <spp>) into profiles/dram_traffic.json: per kernel the number of
launches, the average DRAM bytes per launch and the share of device time.  bench.py reports `roofline.traffic` from it.
    python tools/dram_traffic.py gpurun_out/traffic_caustic.csv caustic 16"""
import collections
import csv
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
src, workload, spp = sys.argv[1], sys.argv[2], int(sys.argv[3])
rows = list(csv.reader(l for l in open(src) if l.startswith('"')))
hdr = rows[0]
ki, mi, vi, ui, ii = (hdr.index(x) for x in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "ID"))
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}
agg = collections.defaultdict(lambda: {"launches": set(), "dram_bytes": 0.0, "us": 0.0})
for r in rows[1:]:
    name = r[ki].split("(")[0].replace("void ", "").replace("trbk::", "")
    v = float(r[vi].replace(",", "")) * SCALE.get(r[ui], 1)
    a = agg[name]
    a["launches"].add(r[ii])
    if r[mi].startswith("dram__bytes"):
        a["dram_bytes"] += v
    elif r[mi].startswith("gpu__time_duration"):
        a["us"] += v
tot_us = sum(a["us"] for a in agg.values())
out = {"workload": workload, "spp": spp, "source": Path(src).name, "kernels": {}}
for k, a in sorted(agg.items(), key=lambda x: -x[1]["us"]):
    n = len(a["launches"])
    out["kernels"][k] = {"launches": n, "dram_bytes_per_launch": a["dram_bytes"] / n, "dram_bytes_total": a["dram_bytes"],
                         "us_total": a["us"], "time_share": a["us"] / tot_us}
f = ROOT / "profiles" / "dram_traffic.json"
allw = json.loads(f.read_text()) if f.exists() else {}
allw[workload] = out
f.write_text(json.dumps(allw, indent=1))
for k, v in out["kernels"].items():
    print(f"{k:32s} n={v['launches']:4d} share={v['time_share']:.3f} dram/launch={v['dram_bytes_per_launch'] / 1e6:9.2f} MB")
