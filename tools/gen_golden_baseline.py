"""tools/gen_golden_baseline.py -- golden whole-sample vectors on the films and depths BASELINE.json names.

    python tools/gen_golden_baseline.py          (needs oracle/_ref/libtr_ref.so, i.e. `make -C oracle ref`)

For configs[1] (caustic) and configs[2] (hard light): film 1024 x 1024, rrDepth 8 and 16 ("max path length 16"), N whole samples of
the UNMODIFIED reference's BDPTIntegrator::render (bdpt.h:219-241) with a private Sampler(seed) each, on camera rays through
random pixel centres.  Stored: rays, seeds, Li, splat pixels / values, draws.  The float streams are NOT stored (they are
std::mt19937(seed) streams: oracle_port.sampler_stream regenerates them, pinned bit-exact in tests/test_oracle.py).
"""
import importlib
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "oracle", ROOT / "tools"):
    sys.path.insert(0, str(p))
trb = importlib.import_module("bidirectional-path-tracing_b200")
import dev_rays  # noqa: E402
import refh  # noqa: E402
import scene_to_obj  # noqa: E402

GOLD = ROOT / "tests" / "golden"
W = H = 1024
N = 600


def main():
    for name in ("caustic", "hardlight"):
        scene = trb.HostScene.load_npz(GOLD / f"scene_{name}.npz")
        obj = scene_to_obj.export_verified(scene, Path(tempfile.mkdtemp(prefix="trb_g_")) / "scene.obj")
        for rr in (8, 16):
            rh = refh.RefHarness.open_scene(obj, width=W, height=H, spp=16, rr_depth=rr)
            rng = np.random.default_rng(31 * rr + len(name))
            cam = dev_rays.camera_rays(rh.camera_matrices(), (0.0, 0.8, 3.8), 30.0, W, H)
            prays = cam[rng.choice(len(cam), N, replace=False)]
            seeds = rng.integers(1, 2**31 - 1, N).astype(np.int32)
            li, pix, rgb, nd = rh.bdpt_paths(prays, seeds, rr - 1)
            rh.close()
            assert nd.max() <= 10 + 10 * (rr - 1)
            np.savez_compressed(GOLD / f"paths_{name}_1024_rr{rr}.npz", rays=prays, seeds=seeds, li=li, splat_pix=pix, splat_rgb=rgb, draws=nd,
                                width=W, height=H, spp=16, rr_depth=rr)
            print(name, rr, "ok: max draws", nd.max(), "splats", int((pix >= 0).sum()))


if __name__ == "__main__":
    main()
