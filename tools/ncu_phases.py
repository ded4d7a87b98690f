"""tools/ncu_phases.py -- per-PHASE shares of a traversal kernel from an .ncu-rep (needs -lineinfo and --import-source on):
    python tools/ncu_phases.py <report.ncu-rep> [launch-index]
Buckets the per-source-line warp-instruction and stall-sample counts of k_trace_wide by marker comments / line ranges of
csrc/kernels/trace_wide.cuh (resolved at run time from the file itself) and of the helpers in trace.cuh."""
import csv
import io
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
K = ROOT / "bidirectional-path-tracing_b200" / "csrc" / "kernels"


def line_of(path, needle, after=0):
    for i, l in enumerate(path.read_text().splitlines(), 1):
        if i > after and needle in l:
            return i
    raise KeyError(needle)


def main():
    rep = sys.argv[1]
    want = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    w, t = K / "trace_wide.cuh", K / "trace.cuh"
    k0 = line_of(w, "__global__ void __launch_bounds__")
    loop = line_of(w, "while (true) {", k0)
    node = line_of(w, "// ---- node phase ----", k0)
    tri = line_of(w, "// ---- triangle phases", k0)
    merge = line_of(w, "if (thit) {", tri)
    end = line_of(w, "#ifdef TRB_COUNTERS", merge)
    pl1 = line_of(w, "auto start_ray", k0)
    try:
        pl0 = line_of(w, "auto push_leaves", k0)      # shipped kernel: the queue append is a lambda before start_ray
    except KeyError:
        pl0 = pl1
    try:
        qa = line_of(w, "// ---- queue append", k0)   # convergent variants: a block between the node phase and the triangle phases
    except KeyError:
        qa = None
    ps0 = line_of(w, "auto push = ", k0)
    box0, box1 = line_of(t, "struct BoxF"), line_of(t, "// rayTriangleIntersect")
    tt0 = line_of(t, "TRB_HD bool tri_test")
    tt_div = line_of(t, "float invDet = frcp(det);")
    tt1 = line_of(t, "// One traversal.")
    pre0, pre1 = line_of(t, "TRB_HD RayPre ray_pre"), line_of(t, "// Decides one child box")

    def bucket(f, ln):
        if f == "trace_wide.cuh":
            if ps0 <= ln < pl0: return "stack push/pop"
            if pl0 <= ln < pl1: return "queue append (push_leaves)"
            if pl1 <= ln < loop: return "start_ray / setup"
            if loop <= ln < node: return "refill + finish (idle lanes)"
            if node <= ln < (qa or tri): return "node phase (own lines)"
            if qa and qa <= ln < tri: return "queue append (prefix sum + stores)"
            if tri <= ln < merge: return "triangle phase: fetch entries, shuffles"
            if merge <= ln < end: return "triangle phase: merge + outst"
            return "kernel other"
        if f == "trace.cuh":
            if box0 <= ln < box1: return "box_fast / box_slow"
            if tt0 <= ln < tt_div: return "tri_test: division-free part"
            if tt_div <= ln < tt1: return "tri_test: division, t, ill"
            if pre0 <= ln < pre1: return "ray_pre"
            if ln < box0: return "ld4 / frcp / box_exact helpers"
            return "trace.cuh other"
        return f
    cur, seen, hdr, kidx = "", set(), None, -1
    agg = {}
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            if kidx < 0 or r[1] in seen:
                kidx += 1; seen = set()
            seen.add(r[1]); cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit() and kidx == want and len(r) > 12:
            ie, te, sm = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
            try:
                v = (int(r[ie]), int(r[te]), int(r[sm]))
            except ValueError:
                continue
            a = agg.setdefault(bucket(cur, int(r[0])), [0, 0, 0])
            for k in range(3):
                a[k] += v[k]
    tot = [sum(a[k] for a in agg.values()) for k in range(3)]
    print(f"launch {want}: {tot[0]} warp instructions, {tot[1] / max(tot[0], 1):.2f} active threads per instruction, {tot[2]} stall samples")
    print(f"{'phase':44s} {'inst':>7s} {'thr/inst':>9s} {'samples':>8s}")
    for name, a in sorted(agg.items(), key=lambda x: -x[1][0]):
        print(f"{name:44s} {a[0] / tot[0]:7.3f} {a[1] / max(a[0], 1):9.1f} {a[2] / max(tot[2], 1):8.3f}")


if __name__ == "__main__":
    main()
