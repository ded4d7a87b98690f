"""tools/exp_sah_upper.py -- experiment: what would a binned-SAH upper tree over the reference's leaves save over the collapse of the
reference's own midpoint tree?  Host build of both, scalar wide traversal (trace_wide.cuh), node visits / slab tests / triangle tests per
ray on seeded ray sets; the hits must be identical (any tree over the leaves finds the reference's candidates).
    python tools/exp_sah_upper.py [rays_per_kind] [tess_tris]"""
import ctypes as C
import importlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "oracle", ROOT / "tools"):
    sys.path.insert(0, str(p))
trb = importlib.import_module("bidirectional-path-tracing_b200")
import check_leaf_claim  # noqa: E402
import gen_tess_scene  # noqa: E402

L = check_leaf_claim.lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 30000
tess = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000


def run(fn, scene, d, rays, any_):
    m = len(rays)
    tuv = np.zeros((m, 3), np.float32); tri = np.zeros(m, np.int32); occ = np.zeros(m, np.uint8); cnt = np.zeros(8, np.uint64)
    assert fn(C.byref(d), rays.ctypes.data_as(C.c_void_p), C.c_long(m), int(any_), tuv.ctypes.data_as(C.c_void_p), tri.ctypes.data_as(C.c_void_p),
              occ.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p)) == 0
    return tuv, tri, occ, cnt


scenes = [(nm, trb.HostScene.load_npz(ROOT / "tests/golden" / f"scene_{nm}.npz")) for nm in ("caustic", "hardlight")]
scenes.append((f"tess {tess}", gen_tess_scene.make_scene(tess, seed=1)))
for name, scene in scenes:
    rays = np.ascontiguousarray(check_leaf_claim.rays_for(scene, n, np.random.default_rng(5)), np.float32)
    d = scene.desc()
    for any_ in (False, True):
        a = run(L.hc_trace_wide, scene, d, rays, any_)
        b = run(L.hc_trace_wide_sah, scene, d, rays, any_)
        same = np.array_equal(a[1], b[1]) and np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32)) and np.array_equal(a[2], b[2])
        m = len(rays)
        print(f"{name:14s} {'any    ' if any_ else 'closest'}: collapse of the reference's tree: nodes/ray {a[3][0] / m:6.2f} slab tests {a[3][1] / m:6.2f} tris {a[3][2] / m:6.2f} | "
              f"SAH over the leaves: nodes/ray {b[3][0] / m:6.2f} slab tests {b[3][1] / m:6.2f} tris {b[3][2] / m:6.2f} | hits identical: {same}")
