"""tools/host_check.py -- DEVELOPMENT AID: runs the __host__ build of csrc/kernels/trace.cuh (tools/host_trace_check.cpp)
against the golden hit records of the unmodified reference, on a box without a GPU.
    python tools/host_check.py [n_extra_random_rays]   (extra rays are checked against oracle/_ref when present)"""
import ctypes as C
import importlib
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "oracle", ROOT / "tools"):
    sys.path.insert(0, str(p))
trb = importlib.import_module("bidirectional-path-tracing_b200")
PKG = ROOT / "bidirectional-path-tracing_b200"
so = Path("/tmp/libhc.so")
objs = [str(PKG / "build" / f"{n}.o") for n in ("obj_loader", "scene_build", "camera", "toml_exr", "host_abi")]
subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off", str(ROOT / "tools/host_trace_check.cpp"), *objs,
                "-o", str(so), "-nostdlib++", "/usr/lib/x86_64-linux-gnu/libstdc++.so.6", "-lm"], check=True)
L = C.CDLL(str(so))


def run(scene, rays, any_, mode=0):
    n = len(rays)
    rays = np.ascontiguousarray(rays, np.float32)
    tuv = np.zeros((n, 3), np.float32); tri = np.zeros(n, np.int32); occ = np.zeros(n, np.uint8); cnt = np.zeros(3, np.uint64)
    d = scene.desc()
    rc = L.hc_trace(C.byref(d), rays.ctypes.data_as(C.c_void_p), C.c_long(n), int(any_), tuv.ctypes.data_as(C.c_void_p),
                    tri.ctypes.data_as(C.c_void_p), occ.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p), mode)
    assert rc == 0
    return tuv, tri, occ, cnt


for name in ("caustic", "hardlight"):
    scene = trb.HostScene.load_npz(ROOT / "tests/golden" / f"scene_{name}.npz")
    g = np.load(ROOT / "tests/golden" / f"trace_{name}.npz")
    tuv, tri, _, cnt = run(scene, g["rays"], False)
    hit = (tri >= 0) & (tuv[:, 0] <= g["rays"][:, 7]) & (tuv[:, 0] >= g["rays"][:, 6])
    tri_v = np.where(hit, tri, -1)
    ok_id = np.array_equal(tri_v, g["tri"])
    m = g["hit"] == 1
    ok_t = np.array_equal(tuv[m].view(np.uint32), g["tuv"][m].view(np.uint32))
    _, _, occ, cnt2 = run(scene, g["rays"], True)
    print(f"{name}: ids {'OK' if ok_id else 'MISMATCH'} tuv {'OK' if ok_t else 'MISMATCH'} occ {'OK' if np.array_equal(occ, g['occluded']) else 'MISMATCH'}"
          f" | closest nodes/ray {cnt[0]/len(tri):.1f} tris/ray {cnt[1]/len(tri):.1f} ill {cnt[2]/len(tri):.4f}"
          f" | any nodes/ray {cnt2[0]/len(tri):.1f} tris/ray {cnt2[1]/len(tri):.1f} ill {cnt2[2]/len(tri):.4f}")
    extra = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    if extra:
        import dev_rays, refh, scene_to_obj, tempfile
        rng = np.random.default_rng(7)
        rays = np.concatenate([dev_rays.random_dir_rays(scene, extra, rng), dev_rays.segment_rays(scene, extra, rng)[0]])
        obj = scene_to_obj.export_verified(scene, Path(tempfile.mkdtemp()) / "s.obj")
        rh = refh.RefHarness.open_scene(obj, width=64, height=64, spp=1, rr_depth=8)
        ids, rtuv, _ = rh.intersect(rays); rocc = rh.occluded(rays)
        base = np.concatenate([[0], np.cumsum(np.bincount(scene.tri_shape, minlength=scene.n_shapes))])
        rtri = np.where(ids[:, 0] == 1, base[np.clip(ids[:, 1], 0, None)] + ids[:, 2], -1)
        tuv, tri, _, cnt = run(scene, rays, False)
        hit = (tri >= 0) & (tuv[:, 0] <= rays[:, 7]) & (tuv[:, 0] >= rays[:, 6])
        _, _, occ, _ = run(scene, rays, True)
        mm = ids[:, 0] == 1
        print(f"  extra {len(rays)} rays vs oracle/_ref: id mismatches {(np.where(hit, tri, -1) != rtri).sum()}, "
              f"tuv mismatches {(tuv[mm].view(np.uint32) != rtuv[mm].view(np.uint32)).any(1).sum()}, occ mismatches {(occ != rocc).sum()}, ill {cnt[2]/len(rays):.4f}")
        rh.close()
