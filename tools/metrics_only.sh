# tools/metrics_only.sh TAG -- ncu per-kernel metric passes (tools/kernel_metrics.py) for the three bench workloads
TAG=${1:-r2x}
cd $GRAFT_REPO_ROOT
M=$(python tools/kernel_metrics.py --metrics)
python tools/prof_render.py caustic 16 > gpurun_out/${TAG}_km_plain.log 2>&1 || exit 1
timeout 1500 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/${TAG}_km_caustic.csv python tools/prof_render.py caustic 16 > gpurun_out/${TAG}_km_caustic.log 2>&1
timeout 1500 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/${TAG}_km_tess.csv python tools/prof_render.py tess 16 > gpurun_out/${TAG}_km_tess.log 2>&1
timeout 1500 ncu --metrics $M --clock-control none --csv --log-file gpurun_out/${TAG}_km_hardlight.csv python tools/prof_render.py hardlight 8 > gpurun_out/${TAG}_km_hardlight.log 2>&1
ls -la gpurun_out/${TAG}_km_*.csv
