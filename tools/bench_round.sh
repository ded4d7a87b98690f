cd $GRAFT_REPO_ROOT
python bench.py --steps 5 --warmup 3 2> gpurun_out/r1n_bench_caustic.err | tail -1 > gpurun_out/r1n_bench_caustic.json
python bench.py --workload tess --spp 4 --steps 5 --warmup 3 2> gpurun_out/r1n_bench_tess.err | tail -1 > gpurun_out/r1n_bench_tess.json
python bench.py --workload hardlight --spp 8 --steps 5 --warmup 3 2> gpurun_out/r1n_bench_hard.err | tail -1 > gpurun_out/r1n_bench_hard.json
tail -2 gpurun_out/r1n_bench_hard.err
