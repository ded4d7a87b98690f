"""tools/gen_golden_direct.py -- level-2 pins for DirectIntegrator: whole samples through the UNMODIFIED reference's
DirectIntegrator::render (src/integrators/direct.h:449-462, `type = "direct"`) with private Samplers, one file per scene and sampling
strategy, plus the sphere-light scene the integrator is written for (tools/gen_tess_scene.make_sphere_light_scene) and the
reference's shape centres / radii (Scene::getShapeCenter / getShapeRadius).

    tests/golden/scene_spherelight.npz, host_spherelight.npz
    tests/golden/dipaths_<scene>_<strategy>.npz   camera rays + Sampler seeds -> Li, draws, settings

Random streams are regenerated from the seeds by the tests (oracle_port.sampler_stream).  Needs /root/reference."""
from __future__ import annotations

import importlib
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "oracle"))
sys.path.insert(0, str(ROOT / "tools"))
import dev_rays  # noqa: E402
import gen_golden  # noqa: E402
import gen_tess_scene  # noqa: E402
import refh  # noqa: E402
import scene_to_obj  # noqa: E402

trb = importlib.import_module("bidirectional-path-tracing_b200")
GOLD = ROOT / "tests" / "golden"
W, H = gen_golden.PATH_W, gen_golden.PATH_H
N = 1500
STRATEGIES = {  # name -> (emitterSamples, bsdfSamples)
    "mis": (2, 2), "area": (3, 1), "solidAngle": (3, 1), "cosineHemisphere": (3, 1), "bsdf": (1, 3),
}


def main():
    sl = gen_tess_scene.make_sphere_light_scene()
    sl_obj = scene_to_obj.export_verified(sl, Path(tempfile.mkdtemp(prefix="trb_sl_")) / "spherelight.obj")
    sl.save_npz(GOLD / "scene_spherelight.npz")
    todo = {"spherelight": sl_obj, "caustic": gen_golden.SCENES["caustic"][0], "tess": gen_golden._tess_obj()}
    for name, obj in todo.items():
        for sname, (m, nb) in STRATEGIES.items():
            extra = f'emitterSamples = {m}\nbsdfSamples = {nb}\nsamplingStrategy = "{sname}"\n'
            rh = refh.RefHarness.open_scene(obj, width=W, height=H, spp=4, rr_depth=5, type_="direct", extra=extra)
            if name == "spherelight" and sname == "mis":
                hs = trb.HostScene.load_obj(obj)
                d = rh.scene_dump()
                for k in ("positions", "normals", "tri_v", "tri_n", "tri_mat", "tri_shape"):
                    assert np.array_equal(getattr(hs, k), d[k]), f"loader mismatch in {k}"
                em = rh.emitters()
                np.savez_compressed(GOLD / "host_spherelight.npz", order=rh.bvh_order(), em_shape=em["shape"], em_area=em["area"],
                                    em_radiance=em["radiance"], em_cdf=np.concatenate(em["cdf"]), camera=rh.camera_matrices(),
                                    width=W, height=H, spheres=rh.shape_spheres(rh.n_shapes))
            rng = np.random.default_rng(20261107 + len(sname))
            cam = dev_rays.camera_rays(rh.camera_matrices(), (0.0, 0.8, 3.8), 30.0, W, H)
            prays = cam[rng.choice(len(cam), N, replace=True)]
            seeds = rng.integers(1, 2**31 - 1, N).astype(np.int32)
            li, nd = rh.integrator_paths(prays, seeds)
            np.savez_compressed(GOLD / f"dipaths_{name}_{sname}.npz", rays=prays, seeds=seeds, li=li, draws=nd, width=W, height=H, spp=4,
                                rr_depth=5, strategy=np.asarray(sname), emitter_samples=m, bsdf_samples=nb,
                                spheres=rh.shape_spheres(rh.n_shapes))
            rh.close()
            print(name, sname, "paths", N, "draws max/mean", nd.max(), nd.mean(), "mean Li", li.mean(), "nonzero", (li.sum(1) > 0).mean(),
                  "nan", np.isnan(li).any())


if __name__ == "__main__":
    main()
