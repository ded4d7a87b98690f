"""tools/prof_multi.py -- the library's own multi-GPU call from ONE process (trb_render_bdpt_multi: host thread per device,
sample-index shards, one ncclReduce): python tools/prof_multi.py [n_gpus] [spp_per_gpu] [width height]"""
import importlib
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
trb = importlib.import_module("bidirectional-path-tracing_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
w, h = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (1024, 1024)
scene = trb.HostScene.load_npz(ROOT / "tests" / "golden" / "scene_caustic.npz")
ctxs = [trb.Context(g) for g in range(n)]
for c in ctxs:
    c.upload(scene)
for k in (1, n):
    d = trb.render_desc(width=w, height=h, spp=spp * k, rr_depth=8)
    for _ in range(2):
        t0 = time.perf_counter()
        img, st = trb.render_bdpt_multi(ctxs[:k], d)
        dt = time.perf_counter() - t0
    print(f"{k} GPU(s): {w}x{h}, {spp * k} spp: {dt * 1e3:.1f} ms wall (host frame buffer out) -> {w * h * spp * k / dt / 1e6:.1f} Msamples/s; mean {img.mean():.4f}")
