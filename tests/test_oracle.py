"""CPU tests: the oracle restatement (oracle/bdpt_oracle.cpp) pinned against golden vectors generated from the
UNMODIFIED reference (tools/gen_golden.py -> tests/golden/*.npz), and the product's host-side logic against the
same vectors.  No GPU needed."""
import numpy as np
import pytest

import oracle_port


@pytest.fixture(scope="module")
def orc(trb, scene_name):
    from conftest import GOLD
    o = oracle_port.Oracle(trb.HostScene.load_npz(GOLD / f"scene_{scene_name}.npz"))
    yield o
    o.close()


def test_sampler_stream_bit_exact(gold):
    g = gold("paths", "caustic")
    for i in (0, 1, 17, 999):
        s = oracle_port.sampler_stream(int(g["seeds"][i]), g["rng"].shape[1])
        assert np.array_equal(s.view(np.uint32), g["rng"][i].view(np.uint32))


def test_bvh_order_emitters_camera(trb, orc, gold, scene_name):
    from conftest import GOLD
    g = gold("host", scene_name)
    hs = trb.HostScene.load_npz(GOLD / f"scene_{scene_name}.npz")
    # oracle restatement of BVH::build and the product's host builder both reproduce the reference's order
    assert np.array_equal(orc.bvh_order(), g["order"])
    assert np.array_equal(hs.bvh_order(), g["order"])
    for em in (orc.emitters(), hs.emitters()):
        assert em["n"] == len(g["em_shape"]) and np.array_equal(em["shape"], g["em_shape"])
        assert np.array_equal(em["area"].view(np.uint32), g["em_area"].view(np.uint32))
        assert np.array_equal(em["cdf"][: len(g["em_cdf"])].view(np.uint32), g["em_cdf"].view(np.uint32))
    desc = trb.render_desc(width=int(g["width"]), height=int(g["height"]))
    orc.set_config(desc)
    assert np.array_equal(orc.camera_matrices().view(np.uint32), g["camera"].view(np.uint32))
    assert np.array_equal(trb.camera_matrices(desc).view(np.uint32), g["camera"].view(np.uint32))


def test_intersection_bit_exact(orc, gold, scene_name):
    g = gold("trace", scene_name)
    ids, tuv = orc.intersect(g["rays"])
    assert np.array_equal(ids[:, 0], g["hit"].astype(np.int32))
    assert np.array_equal(ids[:, 1], g["shape"]) and np.array_equal(ids[:, 2], g["prim"])
    assert np.array_equal(tuv.view(np.uint32), g["tuv"].view(np.uint32))
    assert np.array_equal(orc.occluded(g["rays"]), g["occluded"])


def test_bsdf_bit_exact(orc, gold, scene_name):
    g = gold("bsdf", scene_name)
    for m in np.unique(g["mat"]):
        sel = g["mat"] == m
        out = orc.bsdf(int(m), g["q"][sel])
        assert np.array_equal(out.view(np.uint32), g["out"][sel].view(np.uint32)), f"material {m}"


PT_CONFIGS = ["rr", "rr95", "mis", "implicit"]


def _pt_settings(g):
    return dict(is_explicit=bool(g["is_explicit"]), max_depth=int(g["max_depth"]), rr_prob=float(g["rr_prob"]),
                emitter_samples=int(g["emitter_samples"]), bsdf_samples=int(g["bsdf_samples"]))


@pytest.mark.parametrize("config", PT_CONFIGS)
def test_path_tracer_bit_exact(trb, orc, gold, scene_name, config):
    """the restatement of PathTracerIntegrator (path.h) against whole samples of the unmodified reference
    (tools/gen_golden_path.py): Li and the number of draws, bit for bit"""
    g = gold("ptpaths", f"{scene_name}_{config}")
    desc = trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=int(g["rr_depth"]))
    orc.set_config(desc)
    orc.set_path(depth_cap=0, **_pt_settings(g))
    li, nd = orc.path_paths(g["rays"], seeds=g["seeds"])
    assert np.array_equal(nd, g["draws"])
    assert np.array_equal(li.view(np.uint32), g["li"].view(np.uint32))
    assert (li > 0).any()


def test_bdpt_paths_russian_roulette_bit_exact(trb, orc, gold, scene_name):
    """the restatement's NO_RR 0 branch against the reference built with its own Russian-roulette code enabled
    (oracle/Makefile ref_rr; vectors by tools/gen_golden_rr.py)"""
    g = gold("paths", scene_name + "_rr")
    desc = trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=int(g["rr_depth"]))
    orc.set_config(desc)
    orc.set_rr(True, 0)
    try:
        k = g["splat_pix"].shape[1]
        li, pix, rgb, nd = orc.bdpt_paths(g["rays"], k, seeds=g["seeds"])
    finally:
        orc.set_rr(False)
    assert np.array_equal(nd, g["draws"])
    assert nd.mean() > 45  # walks really run past rr_depth 4 (the capped build draws < 30 per path there)
    assert np.array_equal(li.view(np.uint32), g["li"].view(np.uint32))
    assert np.array_equal(pix, g["splat_pix"])
    assert np.array_equal(rgb.view(np.uint32), g["splat_rgb"].view(np.uint32))


def test_bdpt_paths_bit_exact(trb, orc, gold, scene_name):
    g = gold("paths", scene_name)
    desc = trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=int(g["rr_depth"]))
    orc.set_config(desc)
    k = g["splat_pix"].shape[1]
    li, pix, rgb, nd = orc.bdpt_paths(g["rays"], k, seeds=g["seeds"])
    assert np.array_equal(nd, g["draws"])
    assert np.array_equal(li.view(np.uint32), g["li"].view(np.uint32))
    assert np.array_equal(pix, g["splat_pix"])
    assert np.array_equal(rgb.view(np.uint32), g["splat_rgb"].view(np.uint32))
    # injected-number mode (what the GPU parity entry point consumes) gives the same answer
    li2, pix2, rgb2, nd2 = orc.bdpt_paths(g["rays"][:500], k, injected=g["rng"][:500])
    assert np.array_equal(li2.view(np.uint32), g["li"][:500].view(np.uint32)) and np.array_equal(nd2, g["draws"][:500])


@pytest.mark.parametrize("name,rr_depth", [("caustic", 8), ("caustic", 16), ("hardlight", 8), ("hardlight", 16)])
def test_bdpt_paths_bit_exact_at_baseline_film_and_depth(trb, gold, name, rr_depth):
    """the restatement on BASELINE.json's films and depths (1024 x 1024; rrDepth 8 and 16 = up to 15 vertices per subpath and 225
    connectVertices calls per sample) against whole samples of the unmodified reference (tools/gen_golden_baseline.py)"""
    from conftest import GOLD
    g = gold("paths", f"{name}_1024_rr{rr_depth}")
    o = oracle_port.Oracle(trb.HostScene.load_npz(GOLD / f"scene_{name}.npz"))
    try:
        o.set_config(trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=rr_depth))
        li, pix, rgb, nd = o.bdpt_paths(g["rays"], rr_depth - 1, seeds=g["seeds"])
    finally:
        o.close()
    assert np.array_equal(nd, g["draws"])
    assert np.array_equal(li.view(np.uint32), g["li"].view(np.uint32))
    assert np.array_equal(pix, g["splat_pix"])
    assert np.array_equal(rgb.view(np.uint32), g["splat_rgb"].view(np.uint32))
    assert (pix >= 0).sum() > 1000


DI_SCENES = ["spherelight", "caustic", "tess"]
DI_STRATEGIES = ["mis", "area", "solidAngle", "cosineHemisphere", "bsdf"]


@pytest.fixture(scope="module")
def di_orc(trb):
    from conftest import GOLD
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = oracle_port.Oracle(trb.HostScene.load_npz(GOLD / f"scene_{name}.npz"))
        return cache[name]
    yield get
    for o in cache.values():
        o.close()


@pytest.mark.parametrize("strategy", DI_STRATEGIES)
@pytest.mark.parametrize("di_scene", DI_SCENES)
def test_direct_integrator_bit_exact(trb, di_orc, gold, di_scene, strategy):
    """the restatement of DirectIntegrator (direct.h) against whole samples of the unmodified reference
    (tools/gen_golden_direct.py): Li and the number of draws, bit for bit, for the five sampling strategies; the sphere-light
    scene is the case the integrator is written for, the other two have quad emitters (its cone then is the shape's x-extent)"""
    g = gold("dipaths", f"{di_scene}_{strategy}")
    o = di_orc(di_scene)
    o.set_config(trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=int(g["rr_depth"])))
    o.set_direct(strategy, int(g["emitter_samples"]), int(g["bsdf_samples"]))
    li, nd = o.direct_paths(g["rays"], seeds=g["seeds"])
    assert np.array_equal(nd, g["draws"])
    assert np.array_equal(li.view(np.uint32), g["li"].view(np.uint32))
    assert (li > 0).any()
    # injected-number mode (what the GPU parity entry point consumes)
    stride = int(g["draws"].max()) + 2
    rng = np.stack([oracle_port.sampler_stream(int(s), stride) for s in g["seeds"][:300]])
    li2, nd2 = o.direct_paths(g["rays"][:300], injected=rng)
    assert np.array_equal(li2.view(np.uint32), g["li"][:300].view(np.uint32)) and np.array_equal(nd2, g["draws"][:300])


@pytest.mark.parametrize("di_scene", DI_SCENES)
def test_shape_spheres_bit_exact(trb, di_orc, gold, di_scene):
    """Scene::getShapeCenter / getShapeRadius (renderer.cpp:294-304, 349-358) of every shape"""
    g = gold("dipaths", f"{di_scene}_mis")
    sp = di_orc(di_scene).shape_spheres(len(g["spheres"]))
    assert np.array_equal(sp.view(np.uint32), g["spheres"].view(np.uint32))


def test_bitmap_textures_bit_exact(trb, gold):
    """BitmapTexture3f (core.h:525-585; map_Kd on Diffuse / Phong, map_Kd + map_Ks on Mixture, wrapped and negative
    texture coordinates) in the restatement against the unmodified reference, which read the exported OBJ + MTL + PPM files itself
    (tools/gen_golden_textured.py): whole BDPT, path-tracer and direct-integrator samples, bit for bit"""
    from conftest import GOLD
    hs = trb.HostScene.load_npz(GOLD / "scene_textured.npz")
    assert len(hs.textures) == 4 and (hs.materials["tex"] != 0).sum() == 4
    o = oracle_port.Oracle(hs)
    try:
        g = gold("paths", "textured")
        o.set_config(trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=int(g["rr_depth"])))
        li, pix, rgb, nd = o.bdpt_paths(g["rays"], g["splat_pix"].shape[1], seeds=g["seeds"])
        assert np.array_equal(nd, g["draws"]) and np.array_equal(li.view(np.uint32), g["li"].view(np.uint32))
        assert np.array_equal(pix, g["splat_pix"]) and np.array_equal(rgb.view(np.uint32), g["splat_rgb"].view(np.uint32))
        g = gold("ptpaths", "textured_mis")
        o.set_config(trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=int(g["rr_depth"])))
        o.set_path(depth_cap=0, **_pt_settings(g))
        li, nd = o.path_paths(g["rays"], seeds=g["seeds"])
        assert np.array_equal(nd, g["draws"]) and np.array_equal(li.view(np.uint32), g["li"].view(np.uint32))
        g = gold("dipaths", "textured_mis")
        o.set_direct("mis", int(g["emitter_samples"]), int(g["bsdf_samples"]))
        li, nd = o.direct_paths(g["rays"], seeds=g["seeds"])
        assert np.array_equal(nd, g["draws"]) and np.array_equal(li.view(np.uint32), g["li"].view(np.uint32))
        # the host builder on the textured scene: the reference's BVH order and emitters
        h = gold("host", "textured")
        assert np.array_equal(hs.bvh_order(), h["order"]) and np.array_equal(o.bvh_order(), h["order"])
    finally:
        o.close()
