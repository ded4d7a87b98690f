# tools/prof_logic.sh -- ncu --set full of the logic kernels (bounce / connect) of one 4 M-sample caustic wave; run under gpurun
cd $GRAFT_REPO_ROOT
python tools/prof_render.py caustic 4 > gpurun_out/r1o_prof_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_connect_vertices|k_eye_bounce|k_light_bounce' --launch-skip 1 --launch-count 12 \
    -o gpurun_out/r1o_logic python tools/prof_render.py caustic 4 > gpurun_out/r1o_ncu_logic.log 2>&1
ls -la gpurun_out/r1o_logic.ncu-rep
