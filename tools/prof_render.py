"""tools/prof_render.py -- the smallest program that drives the render path, for ncu (one GPU, short).
    python tools/prof_render.py [workload] [spp] [wave]"""
import importlib
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
trb = importlib.import_module("bidirectional-path-tracing_b200")
wl = sys.argv[1] if len(sys.argv) > 1 else "caustic"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 2
wave = int(sys.argv[3]) if len(sys.argv) > 3 else 0
rr = {"caustic": 8, "hardlight": 16, "tess": 8}[wl]
ctx = trb.Context(0)
if wl == "tess":  # python tools/prof_render.py tess <spp> <wave> <tris>
    sys.path.insert(0, str(ROOT / "tools"))
    import gen_tess_scene
    ctx.upload(gen_tess_scene.make_scene(int(sys.argv[4]) if len(sys.argv) > 4 else 10_000_000, seed=1))
else:
    ctx.upload(trb.HostScene.load_npz(ROOT / "tests" / "golden" / f"scene_{wl}.npz"))
import os
desc = trb.render_desc(width=1024, height=1024, spp=spp, rr_depth=rr, wave_samples=wave, flags=int(os.environ.get("TRB_FLAGS", "0")))
img, st = ctx.render_bdpt(desc)
print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in st.items()}, "mean", float(img.mean()))
