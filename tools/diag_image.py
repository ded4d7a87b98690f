"""tools/diag_image.py -- DEVELOPMENT AID: render a pinned config on the GPU and save the float image (npy) for
offline comparison against the reference's EXRs.  python tools/diag_image.py <scene> <rr_depth> <spp> <out.npy>"""
import importlib, sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
trb = importlib.import_module("bidirectional-path-tracing_b200")
scene, rr, spp, out = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
ctx = trb.Context(0); ctx.upload(trb.HostScene.load_npz(ROOT / "tests/golden" / f"scene_{scene}.npz"))
img, st = ctx.render_bdpt(trb.render_desc(width=800, height=600, spp=spp, rr_depth=rr))
np.save(out, img.astype(np.float16))
print(scene, rr, spp, "mean", img.mean(), "%.1f Msamples/s" % (st["samples"] / st["ms_total"] / 1e3))
