# tools/bench_round2.sh TAG -- the round-2 measurement set on one B200 (run under gpurun)
TAG=${1:-r2x}
cd $GRAFT_REPO_ROOT
python bench.py --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/${TAG}_bench_caustic.err | tail -1 > gpurun_out/${TAG}_bench_caustic.json
python bench.py --workload tess --spp 4 --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/${TAG}_bench_tess.err | tail -1 > gpurun_out/${TAG}_bench_tess.json
python bench.py --workload hardlight --spp 8 --steps 5 --warmup 3 --no-cpu-baseline 2> gpurun_out/${TAG}_bench_hard.err | tail -1 > gpurun_out/${TAG}_bench_hard.json
tail -2 gpurun_out/${TAG}_bench_hard.err
for w in caustic tess hard; do python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_bench_$w.json"))
    print("$w", d["value"], d.get("e2e", {}).get("value"), d.get("stage_ms_serial"), d.get("roofline", {}).get("frac"), d.get("ms_per_step"))
except Exception as e:
    print("$w", "failed", e)
PY
done
