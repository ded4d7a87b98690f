"""tools/gen_golden.py -- regenerate tests/golden/* from the UNMODIFIED reference (oracle/_ref/libtr_ref.so).

Needs /root/reference (scene files) and `make -C oracle ref`.  Outputs, all seeded and reproducible:
  tests/golden/scene_<name>.npz   flat scene arrays from the product's OBJ loader (verified here, array by
                                  array, against what the reference's tinyobj + Scene::load produced)
  tests/golden/trace_<name>.npz   rays -> AcceleratorBVH::intersect / BVH::getIntersection(occlusion) results
  tests/golden/bsdf_<name>.npz    per material (wo, wi, u) -> BSDF::eval/pdf/sample results
  tests/golden/paths_<name>.npz   camera rays + per-path Sampler seeds -> BDPTIntegrator::render Li and splats,
                                  plus the Sampler float streams themselves
  tests/golden/host_<name>.npz    BVH primitive order, emitters, camera matrices
"""
import importlib
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "oracle")); sys.path.insert(0, str(ROOT / "tools"))
import dev_rays  # noqa: E402
import refh  # noqa: E402

trb = importlib.import_module("bidirectional-path-tracing_b200")
REF = Path(os.environ.get("TRB_REFERENCE", "/root/reference"))
GOLD = ROOT / "tests" / "golden"
SCENES = {
    "caustic": (REF / "data/a5/cbox/mesh/cbox_mirror.obj", 8),     # glass sphere, configs C1/C2/C4
    "hardlight": (REF / "data/a5/bonus_bdpt/mesh/cbox.obj", 8),    # occluded emitter, configs C1/C3
}


def _tess_obj():
    """the synthetic scene of configs[4] at 20 k triangles (tools/gen_tess_scene.py): brings PhongBSDF (illum 2) and
    MirrorBSDF (illum 3), which no shipped scene uses, under the reference's own pins"""
    import tempfile
    import gen_tess_scene
    import scene_to_obj
    return scene_to_obj.export_verified(gen_tess_scene.make_scene(20_000, seed=1), Path(tempfile.mkdtemp(prefix="trb_tess_")) / "tess.obj")


PATH_W, PATH_H = 96, 72  # small film so the splat scan of the harness stays cheap
N_TRACE, N_BSDF, N_PATHS, STRIDE = 8000, 1500, 3000, 128


def main():
    GOLD.mkdir(parents=True, exist_ok=True)
    todo = dict(SCENES)
    if "--only-tess" in sys.argv:
        todo = {}
    if "--tess" in sys.argv or "--only-tess" in sys.argv:
        todo["tess"] = (_tess_obj(), 8)
    for name, (obj, rr) in todo.items():
        hs = trb.HostScene.load_obj(obj)
        rh = refh.RefHarness.open_scene(obj, width=PATH_W, height=PATH_H, spp=4, rr_depth=rr)
        d = rh.scene_dump()
        for k in ("positions", "normals", "tri_v", "tri_n", "tri_mat", "tri_shape"):
            assert np.array_equal(getattr(hs, k), d[k]), f"loader mismatch in {k}"
        hs.save_npz(GOLD / f"scene_{name}.npz")
        em = rh.emitters()
        np.savez_compressed(GOLD / f"host_{name}.npz", order=rh.bvh_order(), em_shape=em["shape"], em_area=em["area"],
                            em_radiance=em["radiance"], em_cdf=np.concatenate(em["cdf"]), camera=rh.camera_matrices(),
                            width=PATH_W, height=PATH_H)
        rng = np.random.default_rng(20261104)
        base = np.concatenate([[0], np.cumsum(np.bincount(hs.tri_shape, minlength=hs.n_shapes))])
        # ---- level 1: intersection ----
        cam = dev_rays.camera_rays(rh.camera_matrices(), (0.0, 0.8, 3.8), 30.0, 128, 96)
        cam = cam[rng.choice(len(cam), N_TRACE, replace=False)]
        rays = np.concatenate([cam, dev_rays.random_dir_rays(hs, N_TRACE, rng), dev_rays.segment_rays(hs, N_TRACE, rng)[0],
                               dev_rays.axis_rays(hs, N_TRACE // 8, rng)])
        ids, tuv, geo = rh.intersect(rays)
        occ = rh.occluded(rays)
        tri = np.where(ids[:, 0] == 1, base[np.clip(ids[:, 1], 0, None)] + ids[:, 2], -1).astype(np.int32)
        np.savez_compressed(GOLD / f"trace_{name}.npz", rays=rays, hit=ids[:, 0].astype(np.uint8), shape=ids[:, 1], prim=ids[:, 2],
                            tri=tri, tuv=tuv, occluded=occ)
        # ---- level 2a: BSDFs ----
        qs, outs, mats = [], [], []
        for m in range(rh.n_materials):
            q = np.zeros((N_BSDF, 8), np.float32)
            for c0 in (0, 3):
                z = rng.random(N_BSDF) * 2 - 1
                ph = rng.random(N_BSDF) * 2 * np.pi
                s = np.sqrt(1 - z * z)
                q[:, c0:c0 + 3] = np.stack([s * np.cos(ph), s * np.sin(ph), z], 1)
            q[:, 6:8] = rng.random((N_BSDF, 2))
            q[: N_BSDF // 10, 6:8] = np.round(q[: N_BSDF // 10, 6:8] * 4) / 4 % 1.0  # exact 0, .25, .5, .75 corner cases
            qs.append(q); outs.append(rh.bsdf(m, q)); mats.append(np.full(N_BSDF, m, np.int32))
        np.savez_compressed(GOLD / f"bsdf_{name}.npz", q=np.concatenate(qs), out=np.concatenate(outs), mat=np.concatenate(mats))
        # ---- level 2b: whole BDPT samples ----
        cam = dev_rays.camera_rays(rh.camera_matrices(), (0.0, 0.8, 3.8), 30.0, PATH_W, PATH_H)
        pick = rng.choice(len(cam), N_PATHS, replace=True)
        prays = cam[pick]
        seeds = rng.integers(1, 2**31 - 1, N_PATHS).astype(np.int32)
        li, pix, rgb, nd = rh.bdpt_paths(prays, seeds, rr - 1)
        assert nd.max() <= STRIDE, nd.max()
        streams = np.stack([refh.sampler_stream(int(s), STRIDE) for s in seeds])
        np.savez_compressed(GOLD / f"paths_{name}.npz", rays=prays, seeds=seeds, rng=streams, li=li, splat_pix=pix, splat_rgb=rgb,
                            draws=nd, width=PATH_W, height=PATH_H, spp=4, rr_depth=rr)
        rh.close()
        print(name, "ok:", hs.n_tris, "tris;", len(rays), "rays;", N_PATHS, "paths; max draws", nd.max())


if __name__ == "__main__":
    main()
