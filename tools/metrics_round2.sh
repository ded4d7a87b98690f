# tools/metrics_round2.sh TAG -- default bench line (with c3 / c5 blocks), the reference arm, and the ncu per-kernel metric passes
TAG=${1:-r2x}
cd $GRAFT_REPO_ROOT
python bench.py --steps 5 --warmup 3 2> gpurun_out/${TAG}_bench_default.err | tail -1 > gpurun_out/${TAG}_bench_default.json
python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 > gpurun_out/${TAG}_bench_ref.json
M=$(python tools/kernel_metrics.py --metrics)
python tools/prof_render.py caustic 16 > gpurun_out/${TAG}_km_plain.log 2>&1 || exit 1
timeout 1500 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/${TAG}_km_caustic.csv python tools/prof_render.py caustic 16 > gpurun_out/${TAG}_km_caustic.log 2>&1
timeout 1500 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/${TAG}_km_tess.csv python tools/prof_render.py tess 16 > gpurun_out/${TAG}_km_tess.log 2>&1
timeout 1500 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/${TAG}_km_hardlight.csv python tools/prof_render.py hardlight 8 > gpurun_out/${TAG}_km_hardlight.log 2>&1
python - <<PY
import json
d = json.load(open("gpurun_out/${TAG}_bench_default.json"))
print("C2", round(d["value"], 1), "e2e", round(d["e2e"]["value"], 1), d["stage_ms_serial"], {k: d["roofline"].get(k) for k in ("bound", "frac", "nodes_per_ray", "tris_per_ray")})
for k in ("c3", "c5"):
    b = d.get(k, {})
    print(k, b.get("value"), b.get("e2e", {}).get("value"), b.get("stage_ms_serial"), {x: b.get("roofline", {}).get(x) for x in ("bound", "frac", "nodes_per_ray")}, b.get("error"))
print("cpu", d.get("cpu_baseline", {}).get("value"), "equal_time", {k: d.get("equal_time", {}).get(k) for k in ("relmse_cpu", "relmse_gpu")})
r = json.load(open("gpurun_out/${TAG}_bench_ref.json")); print("ref arm", r.get("value"), r.get("native_libraries"))
PY
