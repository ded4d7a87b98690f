"""tools/gen_golden_rr.py -- level-2 pins for TRB_FLAG_RUSSIAN_ROULETTE.

The reference ships with `#define NO_RR 1` (src/integrators/bdpt.h:18); its Russian-roulette code is there but
compiled out.  `make -C oracle ref_rr` builds the same sources with that one line flipped (oracle/_ref/libtr_ref_rr.so);
this script runs whole BDPT samples through it, exactly as tools/gen_golden.py does for the shipped build:

    tests/golden/paths_<name>_rr.npz   camera rays + per-path Sampler seeds -> Li, splats, draws (rr_depth 4)

The random streams are not stored (a long walk draws > 300 numbers): tests regenerate them from the seeds with the
oracle's std::mt19937 + uniform_real_distribution<float> (oracle_port.sampler_stream, itself pinned in test_oracle).
Needs /root/reference.  Run: python tools/gen_golden_rr.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "oracle"))
sys.path.insert(0, str(ROOT / "tools"))
import dev_rays  # noqa: E402
import gen_golden  # noqa: E402
import refh  # noqa: E402

GOLD = ROOT / "tests" / "golden"
N_PATHS, RR_DEPTH, MAX_SPLATS = 1500, 4, 63
W, H = gen_golden.PATH_W, gen_golden.PATH_H


def main():
    todo = {k: v[0] for k, v in gen_golden.SCENES.items()}
    todo["tess"] = gen_golden._tess_obj()
    for name, obj in todo.items():
        rh = refh.RefHarness.open_scene(obj, rr=True, width=W, height=H, spp=4, rr_depth=RR_DEPTH)
        rng = np.random.default_rng(20261105)
        cam = dev_rays.camera_rays(rh.camera_matrices(), (0.0, 0.8, 3.8), 30.0, W, H)
        prays = cam[rng.choice(len(cam), N_PATHS, replace=True)]
        seeds = rng.integers(1, 2**31 - 1, N_PATHS).astype(np.int32)
        li, pix, rgb, nd = rh.bdpt_paths(prays, seeds, MAX_SPLATS)
        nsp = (pix >= 0).sum(1)
        keep = int(nsp.max())
        np.savez_compressed(GOLD / f"paths_{name}_rr.npz", rays=prays, seeds=seeds, li=li, splat_pix=pix[:, :keep], splat_rgb=rgb[:, :keep],
                            draws=nd, width=W, height=H, spp=4, rr_depth=RR_DEPTH)
        rh.close()
        print(name, "rr paths:", N_PATHS, "max draws", nd.max(), "mean draws", nd.mean(), "max splats", keep)


if __name__ == "__main__":
    main()
