cd $GRAFT_REPO_ROOT
for w in tess:32 caustic:32; do
  n=${w%%:*}; spp=${w##*:}
  for wave in 16777216 33554432; do
    python bench.py --workload $n --spp $spp --wave $wave --steps 3 --warmup 2 --no-cpu-baseline --no-visits --no-extras 2> gpurun_out/wave.err | tail -1 > gpurun_out/wave.json
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/wave.json"))
    print("$n $spp spp wave $wave:", round(d["value"], 1), "Msamples/s,", round(d["ms_per_step"], 1), "ms/step, serial", {k: round(v, 1) for k, v in d["stage_ms_serial"].items()})
except Exception as e:
    print("$n wave $wave failed", e); print(open("gpurun_out/wave.err").read()[-400:])
PY
  done
done
