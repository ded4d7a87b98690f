# tools/prof_logic.sh TAG -- ncu --set full of the logic kernels (bounce / connect) of one 16 M-sample caustic wave; run under gpurun
TAG=${1:-r8c}
cd $GRAFT_REPO_ROOT
python tools/prof_render.py caustic 16 > gpurun_out/${TAG}_prof_plain.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_connect_vertices|k_eye_bounce|k_light_bounce' --launch-skip 2 --launch-count 12 \
    -o gpurun_out/${TAG}_logic -f python tools/prof_render.py caustic 16 > gpurun_out/${TAG}_ncu_logic.log 2>&1
ls -la gpurun_out/${TAG}_logic.ncu-rep
