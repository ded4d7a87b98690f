# tools/prof_wide.sh TAG -- ncu --set full of two launches of k_trace_wide (closest, any) on the caustic and tessellated scenes
TAG=${1:-r2x}
cd $GRAFT_REPO_ROOT
python tools/prof_render.py caustic 4 > gpurun_out/${TAG}_prof_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_trace_wide --launch-skip 16 --launch-count 2 -o gpurun_out/${TAG}_wide_caustic -f python tools/prof_render.py caustic 4 > gpurun_out/${TAG}_ncu_caustic.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_trace_wide --launch-skip 16 --launch-count 2 -o gpurun_out/${TAG}_wide_tess -f python tools/prof_render.py tess 4 > gpurun_out/${TAG}_ncu_tess.log 2>&1
ls -la gpurun_out/${TAG}_*.ncu-rep
