"""tools/gen_golden_path.py -- level-2 pins for the path tracer: whole samples through the UNMODIFIED reference's
PathTracerIntegrator::render (src/integrators/path.h) with private Samplers, as tools/gen_golden.py does for BDPT.

    tests/golden/ptpaths_<scene>_<config>.npz   camera rays + Sampler seeds -> Li, draws, and the [renderer] settings used

Configs: `rr` = the shipped cbox_bdpt_path.toml settings (explicit, maxDepth -1, rrDepth 5, 1 emitter sample) with rrProb 0.8 so
that nearly all walks stay under the device's 64-vertex cap; `rr95` = the same at the shipped rrProb 0.95 (restatement pin);
`mis` = explicit, maxDepth 3, 2 emitter + 2 BSDF samples (both MIS branches); `implicit` = maxDepth 4.
Random streams are regenerated from the seeds by the tests (oracle_port.sampler_stream).  Needs /root/reference."""
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
W, H = gen_golden.PATH_W, gen_golden.PATH_H
CONFIGS = {  # name -> (n paths, rr_depth, settings)
    "rr": (1200, 5, dict(is_explicit=True, max_depth=-1, rr_prob=0.8, emitter_samples=1, bsdf_samples=0)),
    "rr95": (300, 5, dict(is_explicit=True, max_depth=-1, rr_prob=0.95, emitter_samples=1, bsdf_samples=0)),
    "mis": (1200, 5, dict(is_explicit=True, max_depth=3, rr_prob=0.95, emitter_samples=2, bsdf_samples=2)),
    "implicit": (1200, 5, dict(is_explicit=False, max_depth=4, rr_prob=0.95, emitter_samples=1, bsdf_samples=0)),
}


def renderer_lines(c):
    return (f"isExplicit = {'true' if c['is_explicit'] else 'false'}\nmaxDepth = {c['max_depth']}\n"
            f"emitterSamples = {c['emitter_samples']}\nbsdfSamples = {c['bsdf_samples']}\n")


def main():
    todo = {k: v[0] for k, v in gen_golden.SCENES.items()}
    todo["tess"] = gen_golden._tess_obj()
    for name, obj in todo.items():
        for cname, (n, rr_depth, c) in CONFIGS.items():
            rh = refh.RefHarness.open_scene(obj, width=W, height=H, spp=4, rr_depth=rr_depth, type_="path", rr_prob=c["rr_prob"],
                                            extra=renderer_lines(c))
            rng = np.random.default_rng(20261106 + len(cname))
            cam = dev_rays.camera_rays(rh.camera_matrices(), (0.0, 0.8, 3.8), 30.0, W, H)
            prays = cam[rng.choice(len(cam), n, replace=True)]
            seeds = rng.integers(1, 2**31 - 1, n).astype(np.int32)
            li, nd = rh.integrator_paths(prays, seeds)
            np.savez_compressed(GOLD / f"ptpaths_{name}_{cname}.npz", rays=prays, seeds=seeds, li=li, draws=nd, width=W, height=H, spp=4,
                                rr_depth=rr_depth, **{k: np.asarray(v) for k, v in c.items()})
            rh.close()
            print(name, cname, "paths", n, "draws max/mean", nd.max(), nd.mean(), "mean Li", li.mean())


if __name__ == "__main__":
    main()
