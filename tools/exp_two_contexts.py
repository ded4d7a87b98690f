"""tools/exp_two_contexts.py -- experiment: does running TWO waves at once (two contexts on one device, each with its own pair of
streams and wave buffers, half the sample range each) fill the tails of the persistent traversal kernels?
    python tools/exp_two_contexts.py [workload] [spp] [n_contexts]"""
import importlib
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
trb = importlib.import_module("bidirectional-path-tracing_b200")
wl = sys.argv[1] if len(sys.argv) > 1 else "caustic"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
nctx = int(sys.argv[3]) if len(sys.argv) > 3 else 2
rr = {"caustic": 8, "hardlight": 16}[wl]
scene = trb.HostScene.load_npz(ROOT / "tests" / "golden" / f"scene_{wl}.npz")
ctxs = [trb.Context(0) for _ in range(nctx)]
for c in ctxs:
    c.upload(scene)


def run(n):
    t0 = time.perf_counter()
    th = []
    for g in range(n):
        d = trb.render_desc(width=1024, height=1024, spp=spp, rr_depth=rr, spp_begin=spp * g // n, spp_end=spp * (g + 1) // n)
        th.append(threading.Thread(target=lambda c=ctxs[g], d=d: c.render_bdpt(d, download=False)))
    for t in th:
        t.start()
    for t in th:
        t.join()
    return time.perf_counter() - t0


for n in (1, nctx, 1, nctx):
    run(n)
    ts = [run(n) for _ in range(5)]
    print(f"{wl} {spp} spp, {n} context(s): {np.median(ts) * 1e3:.1f} ms  -> {1024 * 1024 * spp / np.median(ts) / 1e6:.1f} Msamples/s", flush=True)
