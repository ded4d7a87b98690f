"""tools/diag_nan.py -- where do non-finite pixels of the DirectIntegrator strategies come from? (development aid)"""
import importlib, sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
trb = importlib.import_module("bidirectional-path-tracing_b200")
ctx = trb.Context(0)
ctx.upload(trb.HostScene.load_npz(ROOT / "tests/golden/scene_spherelight.npz"))
for seed in (0, 1, 2):
    for s in ("mis", "solidAngle", "bsdf", "cosineHemisphere", "area"):
        img, _ = ctx.render_direct(trb.render_desc(width=160, height=120, spp=256, seed=seed), trb.direct_desc(s, 2, 2))
        bad = ~np.isfinite(img).all(2)
        print(seed, s, "mean", float(np.nanmean(img)), "non-finite pixels", int(bad.sum()), np.argwhere(bad)[:3].tolist())
