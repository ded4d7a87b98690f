# tools/sanitize_round2.sh TAG -- the new GPU tests, then compute-sanitizer memcheck and racecheck over tools/sanitize_small.py
TAG=${1:-r2x}
cd $GRAFT_REPO_ROOT
python -m pytest tests/test_gpu_image.py -m gpu -x -q -s -k "patched_reference or cpp_host_api or multi_gpu_entry or cli_and_cpp" > gpurun_out/${TAG}_new_tests.log 2>&1; echo rc=$? >> gpurun_out/${TAG}_new_tests.log
grep -E "patched reference|passed|failed|rc=" gpurun_out/${TAG}_new_tests.log | tail -5
python tools/sanitize_small.py > gpurun_out/${TAG}_sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_sanitize_plain.log; exit 1; }
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_small.py > gpurun_out/${TAG}_memcheck.log 2>&1; echo "memcheck rc=$?" | tee -a gpurun_out/${TAG}_memcheck.log
grep -E "ERROR SUMMARY|Invalid|memcheck rc" gpurun_out/${TAG}_memcheck.log | head -8
timeout 2400 compute-sanitizer --tool racecheck --racecheck-report analysis --error-exitcode 9 python tools/sanitize_small.py > gpurun_out/${TAG}_racecheck.log 2>&1; echo "racecheck rc=$?" | tee -a gpurun_out/${TAG}_racecheck.log
grep -E "RACECHECK SUMMARY|hazard|Race reported|racecheck rc" gpurun_out/${TAG}_racecheck.log | sort | uniq -c | head -20
