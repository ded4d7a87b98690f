# tools/prof_round2.sh TAG -- A/B of the two traversal kernels on one box + ncu source-level captures of k_trace_wide
TAG=${1:-r2x}
cd $GRAFT_REPO_ROOT
for mode in wide binary; do
  for w in caustic tess; do
    spp=16; [ $w = tess ] && spp=4
    TRB_TRAVERSAL=$mode python bench.py --workload $w --spp $spp --steps 4 --warmup 3 --no-cpu-baseline 2> gpurun_out/${TAG}_ab_${mode}_$w.err | tail -1 > gpurun_out/${TAG}_ab_${mode}_$w.json
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_ab_${mode}_$w.json"))
    print("$mode $w", round(d["value"], 1), {k: round(v, 1) for k, v in d["stage_ms_serial"].items()})
except Exception as e:
    print("$mode $w failed", e)
PY
  done
done
python tools/prof_render.py caustic 4 > gpurun_out/${TAG}_prof_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_trace_wide --launch-skip 16 --launch-count 2 -o gpurun_out/${TAG}_wide_caustic -f python tools/prof_render.py caustic 4 > gpurun_out/${TAG}_ncu_caustic.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_trace_wide --launch-skip 16 --launch-count 2 -o gpurun_out/${TAG}_wide_tess -f python tools/prof_render.py tess 4 > gpurun_out/${TAG}_ncu_tess.log 2>&1
ls -la gpurun_out/*.ncu-rep
