cd $GRAFT_REPO_ROOT
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/r1o_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1 >> gpurun_out/r1o_pytest_gpu.log
python bench.py --steps 5 --warmup 3 2> gpurun_out/r1o_bench_caustic.err | tail -1 > gpurun_out/r1o_bench_caustic.json
python bench.py --workload tess --spp 4 --steps 5 --warmup 3 2> gpurun_out/r1o_bench_tess.err | tail -1 > gpurun_out/r1o_bench_tess.json
python bench.py --workload hardlight --steps 3 --warmup 3 2> gpurun_out/r1o_bench_hard.err | tail -1 > gpurun_out/r1o_bench_hard.json
python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -1 > gpurun_out/r1o_bench_ref.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1o_launches_bench_caustic.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/r1o_ncu.log 2>&1
tail -3 gpurun_out/r1o_pytest_gpu.log; cat gpurun_out/r1o_bench_caustic.json | cut -c1-600
