"""tools/gen_golden_textured.py -- pins for bitmap textures (BitmapTexture3f, core.h:525-585): the generated textured
scene (tools/gen_tess_scene.make_textured_scene: map_Kd on Diffuse / Phong, map_Kd + map_Ks on Mixture, wrapping and negative
texture coordinates) through the UNMODIFIED reference, which loads the exported OBJ + MTL + PPM files itself:

    tests/golden/scene_textured.npz, host_textured.npz, trace_textured.npz
    tests/golden/paths_textured.npz             whole BDPT samples (Li, splats, draws)
    tests/golden/ptpaths_textured_mis.npz       PathTracerIntegrator (explicit, maxDepth 3, 2 + 2 samples)
    tests/golden/dipaths_textured_mis.npz       DirectIntegrator (mis, 2 + 2 samples)

Needs /root/reference."""
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
N, RR, STRIDE = 2000, 8, 128


def main():
    sc = gen_tess_scene.make_textured_scene()
    obj = scene_to_obj.export_verified(sc, Path(tempfile.mkdtemp(prefix="trb_tx_")) / "textured.obj")
    sc.save_npz(GOLD / "scene_textured.npz")
    rng = np.random.default_rng(20261108)
    # ---- BDPT ----
    rh = refh.RefHarness.open_scene(obj, width=W, height=H, spp=4, rr_depth=RR)
    d = rh.scene_dump()
    for k in ("positions", "normals", "tri_v", "tri_n", "tri_mat", "tri_shape"):
        assert np.array_equal(getattr(sc, k), d[k]), f"loader mismatch in {k}"
    em = rh.emitters()
    np.savez_compressed(GOLD / "host_textured.npz", order=rh.bvh_order(), em_shape=em["shape"], em_area=em["area"], em_radiance=em["radiance"],
                        em_cdf=np.concatenate(em["cdf"]), camera=rh.camera_matrices(), width=W, height=H)
    cam = dev_rays.camera_rays(rh.camera_matrices(), (0.0, 0.8, 3.8), 30.0, W, H)
    prays = cam[rng.choice(len(cam), N, replace=True)]
    seeds = rng.integers(1, 2**31 - 1, N).astype(np.int32)
    li, pix, rgb, nd = rh.bdpt_paths(prays, seeds, RR - 1)
    assert nd.max() <= STRIDE
    streams = np.stack([refh.sampler_stream(int(s), STRIDE) for s in seeds])
    np.savez_compressed(GOLD / "paths_textured.npz", rays=prays, seeds=seeds, rng=streams, li=li, splat_pix=pix, splat_rgb=rgb, draws=nd,
                        width=W, height=H, spp=4, rr_depth=RR)
    print("bdpt: mean Li", li.mean(), "max draws", nd.max())
    rh.close()
    # ---- path tracer ----
    c = dict(is_explicit=True, max_depth=3, rr_prob=0.95, emitter_samples=2, bsdf_samples=2)
    rh = refh.RefHarness.open_scene(obj, width=W, height=H, spp=4, rr_depth=5, type_="path", rr_prob=0.95,
                                    extra="isExplicit = true\nmaxDepth = 3\nemitterSamples = 2\nbsdfSamples = 2\n")
    li, nd = rh.integrator_paths(prays[:1200], seeds[:1200])
    np.savez_compressed(GOLD / "ptpaths_textured_mis.npz", rays=prays[:1200], seeds=seeds[:1200], li=li, draws=nd, width=W, height=H, spp=4,
                        rr_depth=5, **{k: np.asarray(v) for k, v in c.items()})
    print("path: mean Li", li.mean(), "max draws", nd.max())
    rh.close()
    # ---- direct ----
    rh = refh.RefHarness.open_scene(obj, width=W, height=H, spp=4, rr_depth=5, type_="direct",
                                    extra='emitterSamples = 2\nbsdfSamples = 2\nsamplingStrategy = "mis"\n')
    li, nd = rh.integrator_paths(prays[:1500], seeds[:1500])
    np.savez_compressed(GOLD / "dipaths_textured_mis.npz", rays=prays[:1500], seeds=seeds[:1500], li=li, draws=nd, width=W, height=H, spp=4,
                        rr_depth=5, strategy=np.asarray("mis"), emitter_samples=2, bsdf_samples=2, spheres=rh.shape_spheres(rh.n_shapes))
    print("direct: mean Li", li.mean(), "max draws", nd.max())
    rh.close()


if __name__ == "__main__":
    main()
