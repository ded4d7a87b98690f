# tools/ab_variants.sh TAG v1 v2 ... -- bench caustic + tess with libtrb_<v>.so variants ("default" = libtrb.so, "binary" = binary tree)
TAG=$1; shift
cd $GRAFT_REPO_ROOT
for v in "$@"; do
  for w in ${WORKLOADS:-caustic tess}; do
    spp=16; [ $w = tess ] && spp=4; [ $w = hardlight ] && spp=8
    lib=bidirectional-path-tracing_b200/libtrb_$v.so; mode=wide
    [ $v = default ] && lib=bidirectional-path-tracing_b200/libtrb.so
    [ $v = binary ] && lib=bidirectional-path-tracing_b200/libtrb.so && mode=binary
    TRB_LIB=$GRAFT_REPO_ROOT/$lib TRB_TRAVERSAL=$mode python bench.py --workload $w --spp $spp --steps 3 --warmup 3 --no-cpu-baseline --no-visits --no-extras 2> gpurun_out/${TAG}_ab_${v}_$w.err | tail -1 > gpurun_out/${TAG}_ab_${v}_$w.json
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_ab_${v}_$w.json"))
    print("$v $w", round(d["value"], 1), {k: round(x, 1) for k, x in d["stage_ms_serial"].items()})
except Exception as e:
    print("$v $w failed", e)
PY
  done
done
