"""tools/check_leaf_claim.py -- brute-force check of the claim the 4-wide upper tree rests on: for a ray whose
direction has three finite non-zero components, a reference leaf whose box passes BBox::intersect (bvh.h:33-69) has only passing
ancestors.  Every (ray, leaf) pair of the three fixture scenes is tested: all boxes of the reference's tree with the reference's six
IEEE divisions, no traversal involved.      python tools/check_leaf_claim.py [rays_per_kind] > profiles/r2_leaf_claim.txt"""
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
import dev_rays  # noqa: E402
PKG = ROOT / "bidirectional-path-tracing_b200"


def lib():
    so = Path("/tmp/libhc.so")
    objs = [str(PKG / "build" / f"{n}.o") for n in ("obj_loader", "scene_build", "camera", "toml_exr", "host_abi", "checkpoint_png")]
    subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-ffp-contract=off", str(ROOT / "tools/host_trace_check.cpp"), *objs,
                    "-o", str(so), "-nostdlib++", "/usr/lib/x86_64-linux-gnu/libstdc++.so.6", "-lm", "-lpthread"], check=True)
    return C.CDLL(str(so))


def check(L, scene, rays):
    rays = np.ascontiguousarray(rays, np.float32)
    out = np.zeros(5, np.uint64)
    d = scene.desc()
    assert L.hc_leaf_implies_ancestors(C.byref(d), rays.ctypes.data_as(C.c_void_p), C.c_long(len(rays)), out.ctypes.data_as(C.c_void_p)) == 0
    return [int(x) for x in out]


def rays_for(scene, n, rng):
    cam = dev_rays.camera_rays(trb.camera_matrices(trb.render_desc(width=320, height=240)), (0.0, 0.8, 3.8), 30.0, 320, 240)
    return np.concatenate([cam[rng.choice(len(cam), min(n, len(cam)), replace=False)], dev_rays.random_dir_rays(scene, n, rng),
                           dev_rays.segment_rays(scene, n, rng)[0], dev_rays.axis_rays(scene, n // 8, rng)])


if __name__ == "__main__":
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
    L = lib()
    total = 0
    for name in ("caustic", "hardlight", "tess"):
        scene = trb.HostScene.load_npz(ROOT / "tests/golden" / f"scene_{name}.npz")
        o = check(L, scene, rays_for(scene, n, np.random.default_rng(3)))
        total += o[2]
        print(f"{name}: {o[0]} non-degenerate rays, {o[1]} (ray, passing leaf) pairs, {o[2]} with a failing ancestor | "
              f"{o[3]} degenerate rays (kept on the binary tree), {o[4]} such pairs among them")
    print("claim holds" if total == 0 else "CLAIM VIOLATED")
