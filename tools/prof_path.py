"""tools/prof_path.py -- path-tracer throughput on the wavefront (one GPU): python tools/prof_path.py [scene] [spp]"""
import importlib
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
trb = importlib.import_module("bidirectional-path-tracing_b200")
name = sys.argv[1] if len(sys.argv) > 1 else "caustic"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
ctx = trb.Context(0)
ctx.upload(trb.HostScene.load_npz(ROOT / "tests" / "golden" / f"scene_{name}.npz"))
d = trb.render_desc(width=1024, height=1024, spp=spp, rr_depth=5)
for _ in range(2):
    img, st = ctx.render_path(d, trb.path_desc())
print(name, "path tracer (explicit, rrProb 0.95):", round(st["samples"] / st["ms_total"] / 1e3, 1), "Msamples/s,",
      round(st["rays_closest"] / st["samples"], 1), "rays/sample,", round(st["rays_closest"] / st["ms_total"] / 1e3, 1), "Mrays/s,",
      {k: round(st[k], 1) for k in ("ms_total", "ms_trace_closest", "ms_logic")}, "mean", float(img.mean()))
