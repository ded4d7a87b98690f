"""tools/time_upload.py -- time-to-first-pixel pieces for the 10 M-triangle scene: python tools/time_upload.py [tris]"""
import importlib
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tools"))
import gen_tess_scene  # noqa: E402

trb = importlib.import_module("bidirectional-path-tracing_b200")
tris = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
sc = gen_tess_scene.make_scene(tris, 1)
for th in ("1", ""):
    if th:
        os.environ["TRB_BUILD_THREADS"] = th
    else:
        os.environ.pop("TRB_BUILD_THREADS", None)
    t = time.time(); sc.bvh_order(); b = time.time() - t
    ctx = trb.Context(0)
    t = time.time(); ctx.upload(sc); u = time.time() - t
    t = time.time(); img, st = ctx.render_bdpt(trb.render_desc(width=1024, height=1024, spp=1, rr_depth=8)); r = time.time() - t
    ctx.close()
    print(f"threads {th or 'default'}: reference-order build {b:.2f} s, trb_scene_upload (build + flatten + H2D) {u:.2f} s, first 1 spp frame {r:.2f} s")
