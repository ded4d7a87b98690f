"""tools/host_check_wide.py -- DEVELOPMENT AID: the host build of csrc/kernels/trace_wide.cuh (4-wide upper tree over the reference's
leaves) against the golden hit records and, with a count argument, against the compiled reference on fresh seeded rays - on a box
without a GPU.   python tools/host_check_wide.py [n_extra_rays_per_kind]"""
import ctypes as C
import importlib
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "oracle", ROOT / "tools"):
    sys.path.insert(0, str(p))
trb = importlib.import_module("bidirectional-path-tracing_b200")
import check_leaf_claim  # noqa: E402  (builds /tmp/libhc.so from tools/host_trace_check.cpp)
L = None


def run(scene, desc, rays, any_, wide=True):
    global L
    if L is None:
        L = check_leaf_claim.lib()
    n = len(rays)
    rays = np.ascontiguousarray(rays, np.float32)
    tuv = np.zeros((n, 3), np.float32); tri = np.zeros(n, np.int32); occ = np.zeros(n, np.uint8); cnt = np.zeros(8, np.uint64)
    args = [C.byref(desc), rays.ctypes.data_as(C.c_void_p), C.c_long(n), int(any_), tuv.ctypes.data_as(C.c_void_p),
            tri.ctypes.data_as(C.c_void_p), occ.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p)]
    rc = L.hc_trace_wide(*args) if wide else L.hc_trace(*args, 0)
    assert rc == 0
    return tuv, tri, occ, cnt


def check(name, scene, rays, ref_tri, ref_tuv, ref_occ):
    d = scene.desc()
    n = len(rays)
    tuv, tri, _, cnt = run(scene, d, rays, False)
    hit = (tri >= 0) & (tuv[:, 0] <= rays[:, 7]) & (tuv[:, 0] >= rays[:, 6])
    tri_v = np.where(hit, tri, -1)
    m = ref_tri >= 0
    ok_id = np.array_equal(tri_v, ref_tri)
    ok_t = np.array_equal(tuv[m].view(np.uint32), ref_tuv[m].view(np.uint32))
    _, _, occ, cnt2 = run(scene, d, rays, True)
    ok_occ = np.array_equal(occ, ref_occ)
    _, _, _, b = run(scene, d, rays, False, wide=False)
    _, _, _, b2 = run(scene, d, rays, True, wide=False)
    print(f"{name}: {n} rays ids {'OK' if ok_id else 'MISMATCH ' + str(int((tri_v != ref_tri).sum()))} tuv {'OK' if ok_t else 'MISMATCH'} occ {'OK' if ok_occ else 'MISMATCH ' + str(int((occ != ref_occ).sum()))}"
          f" | wide closest: nodes/ray {cnt[0]/n:.2f} boxes/ray {cnt[1]/n:.2f} tris/ray {cnt[2]/n:.2f} max sp {cnt[3]} | any: nodes {cnt2[0]/n:.2f} boxes {cnt2[1]/n:.2f} tris {cnt2[2]/n:.2f}"
          f" | binary closest: nodes {b[0]/n:.2f} tris {b[1]/n:.2f}; any: nodes {b2[0]/n:.2f} tris {b2[1]/n:.2f} | degenerate {cnt[4]} | wnodes {cnt[5]} bnodes {cnt[6]} depth {cnt[7]}")
    return ok_id and ok_t and ok_occ


def main():
    ok = True
    extra = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    for name in ("caustic", "hardlight", "tess"):
        scene = trb.HostScene.load_npz(ROOT / "tests/golden" / f"scene_{name}.npz")
        g = np.load(ROOT / "tests/golden" / f"trace_{name}.npz")
        ok &= check(name + " golden", scene, g["rays"], g["tri"], g["tuv"], g["occluded"])
        if extra:
            import dev_rays, refh, scene_to_obj
            rng = np.random.default_rng(7)
            cam = dev_rays.camera_rays(trb.camera_matrices(trb.render_desc(width=320, height=240)), (0.0, 0.8, 3.8), 30.0, 320, 240)
            rays = np.concatenate([cam[rng.choice(len(cam), min(extra, len(cam)), replace=False)], dev_rays.random_dir_rays(scene, extra, rng),
                                   dev_rays.segment_rays(scene, extra, rng)[0], dev_rays.axis_rays(scene, extra // 8, rng)])
            obj = scene_to_obj.export_verified(scene, Path(tempfile.mkdtemp()) / "s.obj")
            rh = refh.RefHarness.open_scene(obj, width=64, height=64, spp=1, rr_depth=8)
            ids, rtuv, _ = rh.intersect(rays); rocc = rh.occluded(rays)
            rh.close()
            base = np.concatenate([[0], np.cumsum(np.bincount(scene.tri_shape, minlength=scene.n_shapes))])
            rtri = np.where(ids[:, 0] == 1, base[np.clip(ids[:, 1], 0, None)] + ids[:, 2], -1).astype(np.int32)
            ok &= check(name + " vs reference", scene, rays, rtri, rtuv, rocc)
    print("ALL OK" if ok else "FAILURES")


if __name__ == "__main__":
    main()
