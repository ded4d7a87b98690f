# launch list of the path tracer on the caustic scene (explicit, 4 spp); run under gpurun
cd $GRAFT_REPO_ROOT
python tools/prof_path.py caustic 4 > gpurun_out/r1o_path_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r1o_launches_path_caustic.csv python tools/prof_path.py caustic 4 > gpurun_out/r1o_path_ncu.log 2>&1
tail -1 gpurun_out/r1o_path_plain.log
