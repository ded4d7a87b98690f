"""tools/sanitize_small.py -- a small pass over every device entry point, for compute-sanitizer (memcheck / racecheck):
    compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import importlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
trb = importlib.import_module("bidirectional-path-tracing_b200")
GOLD = ROOT / "tests" / "golden"
for name in ("caustic", "tess"):
    ctx = trb.Context(0)
    # the tessellated fixture goes through the BVH build on the device (k_bb_*, k_wide_*), the Cornell scene through the host build
    ctx.upload(trb.HostScene.load_npz(GOLD / f"scene_{name}.npz"), build="device" if name == "tess" else "host")
    g = np.load(GOLD / f"trace_{name}.npz")
    r8 = g["rays"][:4000]
    r = np.zeros(len(r8), trb.RAY_DTYPE)
    r["o"], r["d"], r["min_t"], r["max_t"] = r8[:, 0:3], r8[:, 3:6], r8[:, 6], r8[:, 7]
    h = ctx.trace_closest(r)
    o = ctx.trace_any(r)
    kw = dict(width=96, height=72, spp=3, rr_depth=8, wave_samples=5000)
    a, st = ctx.render_bdpt(trb.render_desc(**kw))
    b, _ = ctx.render_bdpt(trb.render_desc(flags=trb.TRB_FLAG_RUSSIAN_ROULETTE, **dict(kw, rr_depth=4)))
    c, _ = ctx.render_bdpt(trb.render_desc(flags=trb.TRB_FLAG_NO_OVERLAP, **kw))
    d, _ = ctx.render_path(trb.render_desc(**dict(kw, rr_depth=5)), trb.path_desc())
    e, _ = ctx.render_path(trb.render_desc(**dict(kw, rr_depth=5)), trb.path_desc(True, 3, 0.95, 2, 2))
    f, _ = ctx.render_path(trb.render_desc(**dict(kw, rr_depth=5)), trb.path_desc(False, 4))
    gp = np.load(GOLD / f"paths_{name}.npz")
    pr = np.zeros(200, trb.RAY_DTYPE)
    p8 = gp["rays"][:200]
    pr["o"], pr["d"], pr["min_t"], pr["max_t"] = p8[:, 0:3], p8[:, 3:6], p8[:, 6], p8[:, 7]
    li, pix, rgb, draws = ctx.bdpt_paths(trb.render_desc(width=int(gp["width"]), height=int(gp["height"]), spp=int(gp["spp"]), rr_depth=int(gp["rr_depth"])),
                                         pr, gp["rng"][:200])
    print(name, "hits", int(h["hit"].sum()), "occluded", int(o.sum()), "means", [round(float(x.mean()), 4) for x in (a, b, c, d, e, f)],
          "paths ok", bool(np.isfinite(li).all()))
    for strat in trb.DI_STRATEGIES:
        di, _ = ctx.render_direct(trb.render_desc(**dict(kw, rr_depth=5)), trb.direct_desc(strat, 2, 2))
        assert np.isfinite(di).all(), strat
    ctx.close()
# bitmap textures: every integrator on the textured scene
ctx = trb.Context(0)
ctx.upload(trb.HostScene.load_npz(GOLD / "scene_textured.npz"))
kw = dict(width=96, height=72, spp=3, rr_depth=6, wave_samples=5000)
t1, _ = ctx.render_bdpt(trb.render_desc(**kw))
t2, _ = ctx.render_path(trb.render_desc(**kw), trb.path_desc(True, 3, 0.95, 2, 2))
t3, _ = ctx.render_direct(trb.render_desc(**kw), trb.direct_desc("mis", 2, 2))
print("textured means", [round(float(x.mean()), 4) for x in (t1, t2, t3)])
ctx.close()
