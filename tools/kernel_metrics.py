"""tools/kernel_metrics.py -- turns an ncu CSV over EVERY launch of tools/prof_render.py <workload> <spp> into
profiles/kernel_metrics.json: per kernel the number of launches, the average DRAM bytes and duration per launch, its share of the
device time, and the issue-slot / active-thread / L1-LSU / L2 / local-memory figures (averaged over the launches).  bench.py reads
`roofline.traffic`, `dram_frac`, `issue_slot_frac`, `active_threads_per_inst`, `l1_lsu_busy_frac` from it.

    ncu --metrics $(python tools/kernel_metrics.py --metrics) --clock-control none --csv --log-file gpurun_out/km_caustic.csv \
        python tools/prof_render.py caustic 16
    python tools/kernel_metrics.py gpurun_out/km_caustic.csv caustic 16
"""
import collections
import csv
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
METRICS = {
    "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write", "gpu__time_duration.sum": "us",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_slots_busy_pct",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "threads_per_inst",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed": "lsu_busy_pct",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct", "l1tex__t_sector_hit_rate.pct": "l1_hit_pct",
    "smsp__inst_executed.sum": "warp_inst", "smsp__sass_inst_executed_op_local_ld.sum": "local_ld_inst",
    "smsp__sass_inst_executed_op_local_st.sum": "local_st_inst",
    "smsp__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}


def main():
    if sys.argv[1] == "--metrics":
        print(",".join(METRICS))
        return
    src, workload, spp = sys.argv[1], sys.argv[2], int(sys.argv[3])
    rows = list(csv.reader(l for l in open(src) if l.startswith('"')))
    hdr = rows[0]
    ki, mi, vi, ui, ii = (hdr.index(x) for x in ("Kernel Name", "Metric Name", "Metric Value", "Metric Unit", "ID"))
    agg = collections.defaultdict(lambda: {"launches": set(), "sum": collections.defaultdict(float), "n": collections.defaultdict(int)})
    for r in rows[1:]:
        name = r[ki].split("(")[0].replace("void ", "").replace("trbk::", "")
        key = METRICS.get(r[mi])
        if key is None:
            continue
        try:
            v = float(r[vi].replace(",", "")) * SCALE.get(r[ui], 1)
        except ValueError:
            continue
        a = agg[name]
        a["launches"].add(r[ii]); a["sum"][key] += v; a["n"][key] += 1
    tot_us = sum(a["sum"]["us"] for a in agg.values())
    out = {"workload": workload, "spp": spp, "source": Path(src).name, "kernels": {}}
    for k, a in sorted(agg.items(), key=lambda x: -x[1]["sum"]["us"]):
        n = len(a["launches"])
        s = a["sum"]
        rec = {"launches": n, "dram_bytes_per_launch": (s["dram_read"] + s["dram_write"]) / n, "us_per_launch": s["us"] / n,
               "us_total": s["us"], "time_share": s["us"] / tot_us}
        for key in ("issue_slots_busy_pct", "threads_per_inst", "lsu_busy_pct", "l2_hit_pct", "l1_hit_pct", "achieved_occupancy_pct"):
            if a["n"][key]:
                rec[key] = s[key] / a["n"][key]
        if s["warp_inst"]:
            rec["local_ld_st_share_of_warp_inst"] = (s["local_ld_inst"] + s["local_st_inst"]) / s["warp_inst"]
        out["kernels"][k] = rec
    f = ROOT / "profiles" / "kernel_metrics.json"
    allw = json.loads(f.read_text()) if f.exists() else {}
    allw[workload] = out
    f.write_text(json.dumps(allw, indent=1))
    for k, v in out["kernels"].items():
        print(f"{k:30s} n={v['launches']:4d} share={v['time_share']:.3f} us/launch={v['us_per_launch']:9.1f} dram/launch={v['dram_bytes_per_launch'] / 1e6:9.2f} MB "
              f"issue={v.get('issue_slots_busy_pct', 0):5.1f}% thr/inst={v.get('threads_per_inst', 0):5.1f} lsu={v.get('lsu_busy_pct', 0):5.1f}% "
              f"L2hit={v.get('l2_hit_pct', 0):5.1f}% local={100 * v.get('local_ld_st_share_of_warp_inst', 0):.2f}%")


if __name__ == "__main__":
    main()
