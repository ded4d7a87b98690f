# tools/exp_wave_size.sh -- samples per wavefront: 2^22 (default) against 2^23 and 2^24 (memory is not the constraint on a 180 GB part)
cd $GRAFT_REPO_ROOT
for w in caustic:32 hardlight:16 tess:16; do
  n=${w%%:*}; spp=${w##*:}
  for wave in 4194304 8388608 16777216; do
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
nvidia-smi --query-gpu=memory.used --format=csv,noheader
