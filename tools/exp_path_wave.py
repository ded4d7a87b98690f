"""tools/exp_path_wave.py -- path-tracer and direct-integrator throughput against the samples per wavefront (one GPU)"""
import importlib
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
trb = importlib.import_module("bidirectional-path-tracing_b200")
for name in ("caustic", "hardlight"):
    ctx = trb.Context(0)
    ctx.upload(trb.HostScene.load_npz(ROOT / "tests" / "golden" / f"scene_{name}.npz"))
    for shift in (21, 22, 23):
        d = trb.render_desc(width=1024, height=1024, spp=16, rr_depth=5, wave_samples=1 << shift)
        for _ in range(2):
            img, st = ctx.render_path(d, trb.path_desc())
        print(name, "path 2^%d:" % shift, round(st["samples"] / st["ms_total"] / 1e3, 1), "Msamples/s", {k: round(st[k], 1) for k in ("ms_total", "ms_trace_closest", "ms_logic")}, flush=True)
    for shift in (22, 23, 24):
        d = trb.render_desc(width=1024, height=1024, spp=16, rr_depth=5, wave_samples=1 << shift)
        for _ in range(2):
            img, st = ctx.render_direct(d, trb.direct_desc("mis", 1, 1))
        print(name, "direct(mis) 2^%d:" % shift, round(st["samples"] / st["ms_total"] / 1e3, 1), "Msamples/s", {k: round(st[k], 1) for k in ("ms_total", "ms_trace_closest", "ms_logic")}, flush=True)
    ctx.close()
