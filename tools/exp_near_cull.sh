# tools/exp_near_cull.sh TAG -- near-side box cull (kernels/trace.cuh: near_cull) off / on, serialised stage times of the three bench workloads
TAG=${1:-r5c}
cd $GRAFT_REPO_ROOT
for nc in 0 1 0 1; do
  for w in caustic hardlight tess; do
    spp=16; [ $w = hardlight ] && spp=8
    TRB_NEAR_CULL=$nc timeout 300 python bench.py --workload $w --spp $spp --steps 3 --warmup 3 --no-cpu-baseline --no-extras 2> gpurun_out/${TAG}_nc${nc}_$w.err | tail -1 > gpurun_out/${TAG}_nc${nc}_$w.json
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_nc${nc}_$w.json"))
    r = d["roofline"]
    print("near_cull=$nc $w", round(d["value"], 1), {k: round(x, 1) for k, x in d["stage_ms_serial"].items()}, "any nodes/tris per ray", round(r.get("nodes_per_ray", 0), 2), round(r.get("tris_per_ray", 0), 2))
except Exception as e:
    print("near_cull=$nc $w failed", e)
PY
  done
done
