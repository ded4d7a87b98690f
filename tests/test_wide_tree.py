"""CPU tests of the 4-wide upper tree the traversal kernels walk (csrc/host/scene_build.cpp: build_wide_nodes,
csrc/kernels/trace_wide.cuh): its structure, the claim it rests on, and the host build of the traversal against the golden hit
records of the unmodified reference.  The GPU twin is tests/test_gpu_parity.py (level 1) and tests/test_gpu_build.py."""
import numpy as np
import pytest

from conftest import GOLD


def test_leaf_box_pass_implies_ancestors_pass(trb, scene_name):
    """For every (ray, reference leaf) pair: the leaf box passes BBox::intersect (bvh.h:33-69) => all its ancestors pass, for rays
    with three finite non-zero direction components (the others stay on the reference's binary tree).  Brute force over every box
    of the reference's tree; profiles/r2_leaf_claim.txt holds the same check on 3.7 M rays."""
    import check_leaf_claim
    L = check_leaf_claim.lib()
    scene = trb.HostScene.load_npz(GOLD / f"scene_{scene_name}.npz")
    o = check_leaf_claim.check(L, scene, check_leaf_claim.rays_for(scene, 2500, np.random.default_rng(9)))
    assert o[0] > 6000 and o[1] > 20000
    assert o[2] == 0, f"{o[2]} passing leaves under a failing ancestor"


def test_wide_traversal_host_build_matches_reference_goldens(trb, gold, scene_name):
    """trace_wide.cuh compiled for the host (the scalar statement of the rule k_trace_wide implements) on the golden rays:
    ids, t, u, v bit-exact, occlusion flags exact; and the wide tree visits fewer than half the nodes of the binary one."""
    import host_check_wide
    scene = trb.HostScene.load_npz(GOLD / f"scene_{scene_name}.npz")
    g = gold("trace", scene_name)
    rays = g["rays"]
    d = scene.desc()
    tuv, tri, _, cnt = host_check_wide.run(scene, d, rays, False)
    hit = (tri >= 0) & (tuv[:, 0] <= rays[:, 7]) & (tuv[:, 0] >= rays[:, 6])
    assert np.array_equal(np.where(hit, tri, -1), g["tri"])
    m = g["tri"] >= 0
    assert np.array_equal(tuv[m].view(np.uint32), g["tuv"][m].view(np.uint32))
    _, _, occ, _ = host_check_wide.run(scene, d, rays, True)
    assert np.array_equal(occ, g["occluded"])
    _, _, _, b = host_check_wide.run(scene, d, rays, False, wide=False)
    assert cnt[0] < 0.5 * b[0], (cnt[0], b[0])
    assert cnt[4] > 0   # the golden set holds degenerate rays: they took the binary tree


def test_wide_tree_structure(trb, scene_name):
    """every reference leaf is reachable exactly once, child boxes nest in their parent's slot box, leaves cover the primitive list"""
    import ctypes as C
    import check_leaf_claim
    L = check_leaf_claim.lib()
    scene = trb.HostScene.load_npz(GOLD / f"scene_{scene_name}.npz")
    d = scene.desc()
    n = C.c_size_t(0)
    assert L.hc_wide_nodes(C.byref(d), None, C.byref(n)) == 0
    w = np.zeros(n.value, trb.DEV_NODE4_DTYPE)
    assert L.hc_wide_nodes(C.byref(d), w.ctypes.data_as(C.c_void_p), C.byref(n)) == 0
    hf = w["half"]                                                          # (n, 2) DevNode-shaped halves: children 0,1 and 2,3
    ref = np.stack([hf["ref0"][:, 0], hf["ref1"][:, 0], hf["ref0"][:, 1], hf["ref1"][:, 1]], 1)
    lo = np.stack([hf["c0min"][:, 0], hf["c1min"][:, 0], hf["c0min"][:, 1], hf["c1min"][:, 1]], 1)
    hi = np.stack([hf["c0max"][:, 0], hf["c1max"][:, 0], hf["c0max"][:, 1], hf["c1max"][:, 1]], 1)
    empty = np.int32(-2**31)
    inner = ref[(ref >= 0)]
    assert np.array_equal(np.sort(inner), np.arange(1, len(w)))            # every node but the root has exactly one parent
    assert (np.diff(inner) > 0).all()                                       # level order: (parent order, slot order)
    leaf = ~ref[(ref < 0) & (ref != empty)]
    first, count = leaf >> 3, leaf & 7
    o = np.argsort(first)
    assert first[o][0] == 0 and np.array_equal(first[o][1:], (first[o] + count[o])[:-1]) and (first[o] + count[o])[-1] == scene.n_tris
    assert (count >= 1).all() and (count <= 4).all()
    # a child node's slot boxes lie inside the box its parent holds for it
    pi, ps = np.nonzero(ref >= 0)
    for p_, s_ in list(zip(pi, ps))[:: max(1, len(pi) // 2000)]:
        c = ref[p_, s_]
        used = ref[c] != empty
        assert (lo[c][used] >= lo[p_, s_]).all() and (hi[c][used] <= hi[p_, s_]).all()
    assert (ref[:, :2] != empty).all()                                       # at least two children per node


def test_sah_upper_tree_is_deterministic_across_thread_counts(trb, monkeypatch):
    """build_sah_wide (csrc/host/scene_build.cpp) on one host thread and on all of them gives the same nodes byte for byte (min / max /
    count reductions and a stable partition: nothing depends on the slicing) - what lets a host-built and a device-built scene agree
    record for record (tests/test_gpu_build.py) on boxes with different core counts."""
    import ctypes as C
    import check_leaf_claim
    import gen_tess_scene
    L = check_leaf_claim.lib()
    scene = gen_tess_scene.make_scene(300_000, seed=4)     # large enough for the parallel top of the tree and the subtree tasks
    d = scene.desc()
    got = []
    for threads in ("1", "3", "16"):
        monkeypatch.setenv("TRB_BUILD_THREADS", threads)
        n = C.c_size_t(0)
        assert L.hc_wide_nodes(C.byref(d), None, C.byref(n)) == 0
        w = np.zeros(n.value, trb.DEV_NODE4_DTYPE)
        assert L.hc_wide_nodes(C.byref(d), w.ctypes.data_as(C.c_void_p), C.byref(n)) == 0
        got.append(w.tobytes())
    assert got[0] == got[1] == got[2] and len(got[0]) > 128 * 1000
