# tools/exp_two_waves.sh -- two waves in flight (two contexts on one device) against one, for traversal grids of 8 / 7 / 6 / 5 blocks per SM
cd $GRAFT_REPO_ROOT
for b in 8 7 6 5; do
  echo "== traversal grid: $b blocks per SM"
  TRB_TRACE_GRID_BLOCKS=$b python tools/exp_two_contexts.py caustic 32 2 2>&1 | tail -2
done
echo "== three contexts, 6 blocks per SM"; TRB_TRACE_GRID_BLOCKS=6 python tools/exp_two_contexts.py caustic 48 3 2>&1 | tail -1
