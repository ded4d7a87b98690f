set -x
cd $GRAFT_REPO_ROOT
P=bidirectional-path-tracing_b200
for v in "" top64 top128 top256 b256top512 b512top1024; do
  if [ -z "$v" ]; then export TRB_LIB=$P/libtrb.so; else export TRB_LIB=$P/libtrb_$v.so; fi
  echo "=== variant: ${v:-default}"
  for rep in 1 2; do
    TRB_FLAGS=2 python tools/prof_render.py caustic 16 2>&1 | tail -1
    TRB_FLAGS=0 python tools/prof_render.py caustic 16 2>&1 | tail -1
  done
  TRB_FLAGS=2 python tools/prof_render.py hardlight 8 2>&1 | tail -1
  TRB_FLAGS=0 python tools/prof_render.py hardlight 8 2>&1 | tail -1
done
export TRB_LIB=$P/libtrb_top128.so
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "trace_closest or trace_any or million or degenerate" 2>&1 | tail -3
