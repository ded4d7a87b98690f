"""GPU tests of SURVEY 8(f) N1: BVH::build (externals/bvh.h:147-247) on the device (csrc/kernels/bvh_build.cuh) must produce the
reference's EXACT tree, primitive order and node numbering - i.e. the same device scene, record for record, as the host build
(which tests/test_oracle.py and tests/test_host.py pin to the unmodified reference's order)."""
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = Path(__file__).parent / "golden"


def _same(a, b):
    """record arrays equal field by field; float fields numerically (a -0 / +0 tie of a min / max may keep either zero)"""
    if a.shape != b.shape:
        return False
    return all(np.array_equal(a[k], b[k]) for k in a.dtype.names)


def _compare(trb, scene, label, wide="sah"):
    host, dev = trb.Context(0), trb.Context(0)
    try:
        host.upload(scene, build="host", wide=wide)
        dev.upload(scene, build="device", wide=wide)
        assert dev.built_on_device() and not host.built_on_device()
        for which in ("nodes", "wnodes", "tris", "shade", "tri_uv"):
            a, b = host.download(which), dev.download(which)
            assert _same(a, b), f"{label}: {which} differ ({len(a)} vs {len(b)} records)"
        assert host.scene_bytes() == dev.scene_bytes()
        # and the scene behaves the same: a small render, and a refresh from the staging blob read back from the device
        desc = trb.render_desc(width=64, height=48, spp=2, rr_depth=6)
        a, _ = host.render_bdpt(desc)
        b, _ = dev.render_bdpt(desc)
        dev.refresh_scene()
        c, _ = dev.render_bdpt(desc)
        assert np.allclose(a, b, rtol=2e-5, atol=2e-6) and np.allclose(a, c, rtol=2e-5, atol=2e-6)
        return len(host.download("nodes"))
    finally:
        host.close(); dev.close()


@pytest.mark.parametrize("name", ["caustic", "hardlight", "tess", "spherelight", "textured"])
def test_device_build_equals_host_build(trb, name):
    n = _compare(trb, trb.HostScene.load_npz(GOLD / f"scene_{name}.npz"), name)
    assert n > 0


def test_device_build_at_size_and_edge_cases(trb, gold):
    import gen_tess_scene
    # TRB_UPLOAD_WIDE_COLLAPSE: the upper tree as the collapse of the reference's tree, made on the device by k_wide_* / on the host by build_wide_nodes
    _compare(trb, trb.HostScene.load_npz(GOLD / "scene_caustic.npz"), "caustic, collapsed upper tree", wide="collapse")
    _compare(trb, gen_tess_scene.make_scene(600_000, seed=3), "tess 600k")   # auto mode takes the device build from 65536 triangles on
    ctx = trb.Context(0)
    ctx.upload(gen_tess_scene.make_scene(100_000, seed=2))
    assert ctx.built_on_device()
    ctx.close()
    # tiny scenes: a single leaf (<= 4 triangles), and five triangles (one split) with coincident centroids (median-index fallback)
    sc = trb.HostScene.load_npz(GOLD / "scene_caustic.npz")
    light = np.flatnonzero(sc.materials["ke"][sc.tri_mat].sum(1) > 0)[:2]
    for pick in (light, np.concatenate([light, light, light[:1]])):
        tiny = trb.HostScene(sc.positions, sc.normals, sc.tri_v[pick], sc.tri_n[pick], sc.tri_mat[pick], np.zeros(len(pick), np.int32), sc.materials, n_shapes=1)
        _compare(trb, tiny, f"{len(pick)} triangles")
    # hits through the device-built tree are the reference's
    g = gold("trace", "tess")
    ctx = trb.Context(0)
    ctx.upload(trb.HostScene.load_npz(GOLD / "scene_tess.npz"), build="device")
    r8 = g["rays"]
    rays = np.zeros(len(r8), trb.RAY_DTYPE)
    rays["o"], rays["d"], rays["min_t"], rays["max_t"] = r8[:, :3], r8[:, 3:6], r8[:, 6], r8[:, 7]
    h = ctx.trace_closest(rays)
    assert np.array_equal(h["tri"], g["tri"]) and np.array_equal(h["t"][g["hit"] == 1].view(np.uint32), g["tuv"][g["hit"] == 1, 0].view(np.uint32))
    assert np.array_equal(ctx.trace_any(rays), g["occluded"])
    ctx.close()


def test_checked_build_runs_clean(trb):
    """compute-sanitizer is closed on this pool, so memory safety of the traversal is checked by the kernels themselves: libtrb_checks.so
    (-DTRB_CHECKS: a device assert on every node / triangle / stack / per-warp-queue / deferred-queue index and on the outstanding-test
    counts, csrc/kernels/trace_coop.cuh) runs tools/sanitize_small.py - every entry point and integrator on the Cornell, tessellated
    (device-built BVH) and textured scenes, tiny waves so that queues refill many times.  A failed assert is a CUDA error and a
    non-zero exit."""
    import os
    import subprocess
    import sys
    root = Path(__file__).resolve().parent.parent
    lib = root / "bidirectional-path-tracing_b200" / "libtrb_checks.so"
    if not lib.exists():
        pytest.skip("libtrb_checks.so not built (python bidirectional-path-tracing_b200/build.py --variant checks -DTRB_CHECKS=1)")
    r = subprocess.run([sys.executable, str(root / "tools" / "sanitize_small.py")], env=dict(os.environ, TRB_LIB=str(lib)), capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "Assertion" not in r.stderr and "textured means" in r.stdout
