"""CPU tests of the host side of the drop-in boundary: the C ABI surface, loadTOML / Scene::load / saveEXR
replacements.  No GPU needed."""
import ctypes
import os
import re
import sys
from pathlib import Path

import numpy as np
import pytest

from conftest import GOLD, ROOT

os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")


def test_abi_exports_every_declared_symbol(trb):
    hdr = (ROOT / "include" / "trb.h").read_text()
    names = set(re.findall(r"^\s*(?:int|void|size_t|const char\*)\s+(trb_[a-z_]+)\s*\(", hdr, re.M))
    assert len(names) >= 20, names
    L = trb.lib()
    missing = [n for n in sorted(names) if not hasattr(L, n)]
    assert not missing, f"libtrb.so does not export {missing}"
    assert "sm_100a" in trb.build_info()


def test_no_cpu_fallback(trb):
    """Without a usable device the device entry points refuse to work (there is no CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(trb.TrbError) as e:
        trb.Context(0)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_load_toml_keys_defaults_and_errors(trb, tmp_path):
    p = tmp_path / "scene.toml"
    p.write_text('[input]\nobjfile = "../mesh/a.obj"\n\n[camera]\neye = [ 0.0, 0.8, 3.8 ]\nat = [ 0.0, 0.8, 0.0 ]\nup = [ 0.0, 1.0, 0.0 ]\n'
                 'fov = 30.0\n\n[film]\nwidth = 800\nheight = 600\n\n[renderer]\nrealtime = false\ntype = "bdpt"\nrrDepth = 8\nrrProb = 0.95\nspp = 256')
    d, obj, typ, rt = trb.load_toml(p)
    assert (d.width, d.height, d.spp, d.rr_depth, typ, rt) == (800, 600, 256, 8, "bdpt", False)
    assert list(d.eye) == [0.0, np.float32(0.8), np.float32(3.8)] and d.fov_deg == 30.0
    assert obj == str(tmp_path / "../mesh/a.obj")  # relative to the toml's directory (renderer.cpp:236-241)
    # defaults of loadTOML (main.cpp:30-46,105,112); an integer array does not satisfy get_array_of<double>
    p.write_text('[input]\nobjfile = "/abs/a.obj"\n[camera]\neye = [1, 2, 3]\n[film]\n[renderer]\ntype = "bdpt"\n')
    d, obj, typ, rt = trb.load_toml(p)
    assert (d.width, d.height, d.spp, d.rr_depth) == (768, 576, 1, 5) and obj == "/abs/a.obj"
    assert list(d.eye) == [1.0, 1.0, 0.0] and list(d.up) == [0.0, 1.0, 0.0] and d.fov_deg == 30.0
    # type = "path" (main.cpp:94-102): defaults, then the shipped cbox_bdpt_path.toml settings and the extra keys
    pd = trb.load_toml_path(p)
    assert (pd.is_explicit, pd.max_depth, pd.emitter_samples, pd.bsdf_samples) == (1, -1, 1, 0) and abs(pd.rr_prob - 0.95) < 1e-7
    p.write_text('[input]\nobjfile = "a.obj"\n[camera]\n[film]\n[renderer]\nrealtime = false\nisExplicit = false\ntype = "path"\nmaxDepth = 4\n'
                 'rrDepth = 7\nrrProb = 0.8\nemitterSamples = 2\nbsdfSamples = 3\nspp = 512\n')
    d, obj, typ, rt = trb.load_toml(p)
    pd = trb.load_toml_path(p)
    assert (typ, d.rr_depth, d.spp) == ("path", 7, 512)
    assert (pd.is_explicit, pd.max_depth, pd.emitter_samples, pd.bsdf_samples) == (0, 4, 2, 3) and abs(pd.rr_prob - 0.8) < 1e-7
    # type = "direct" (main.cpp:88-93): bsdfSamples defaults to 1 here (0 for "path"); the default strategy name, "emitter", is
    # one DirectIntegrator::render rejects (direct.h:449-461)
    dd = trb.load_toml_direct(p)
    assert (dd.strategy, dd.emitter_samples, dd.bsdf_samples) == (-1, 2, 3)
    p.write_text('[input]\nobjfile = "a.obj"\n[camera]\n[film]\n[renderer]\ntype = "direct"\nsamplingStrategy = "solidAngle"\nspp = 4\n')
    dd = trb.load_toml_direct(p)
    assert (trb.load_toml(p)[2], dd.strategy, dd.emitter_samples, dd.bsdf_samples) == ("direct", trb.DI_STRATEGIES["solidAngle"], 1, 1)
    for name, code in trb.DI_STRATEGIES.items():
        p.write_text(f'[input]\nobjfile = "a.obj"\n[camera]\n[film]\n[renderer]\ntype = "direct"\nsamplingStrategy = "{name}"\n')
        assert trb.load_toml_direct(p).strategy == code == trb.direct_desc(name).strategy
    p.write_text('[input]\nobjfile = "a.obj"\n[camera]\n[film]\n')
    with pytest.raises(trb.TrbError):
        trb.load_toml(p)
    with pytest.raises(trb.TrbError):
        trb.load_toml(tmp_path / "missing.toml")


OBJ = """mtllib m.mtl
o quad
v 0.2 1.585 0.2
v -0.2 1.585 0.2
v 0.2 1.585 -0.2
v -0.2 1.585 -0.2
vn 0 -1 0
usemtl light
f 1//1 2//1 4//1 3//1
g second
v 0 0 0
v 1 0 0
v 0 1 0
vn 0 0 1
usemtl white
f -3//-1 -2//-1 -1//-1
usemtl grey
f 5//2 7//2 6//2
"""
MTL = "newmtl light\nKd 0.78 0.78 0.78\nKe 35 35 35\nillum 7\n\nnewmtl white\nKd 0.725 0.71 0.68\nKs 0.5 0.5 0.5\nNs 9.803922\nNi 1.5\nTf 1 1 1\nillum 8\n\nnewmtl grey\nKd 0.3 0.3 0.3\nillum 7\n"


def test_obj_loader_semantics(trb, tmp_path):
    (tmp_path / "s.obj").write_text(OBJ)
    (tmp_path / "m.mtl").write_text(MTL)
    s = trb.HostScene.load_obj(tmp_path / "s.obj")
    assert s.n_shapes == 2 and s.shape_names == ["quad", "second"]
    # ear clipping of the quad 1 2 4 3 (tiny_obj_loader.h:1043-1239): (1,2,4) then (1,4,3)
    assert s.tri_v.tolist() == [[0, 1, 3], [0, 3, 2], [4, 5, 6], [4, 6, 5]]
    assert s.tri_n.tolist() == [[0, 0, 0], [0, 0, 0], [1, 1, 1], [1, 1, 1]]
    assert s.tri_mat.tolist() == [0, 0, 1, 2] and s.tri_shape.tolist() == [0, 0, 1, 1]
    assert s.materials["illum"].tolist() == [7, 8, 7] and s.materials["ns"][0] == 1.0 and s.materials["ni"][1] == 1.5
    assert s.positions[0].tolist() == [np.float32(0.2), np.float32(1.585), np.float32(0.2)]
    em = s.emitters()
    assert em["n"] == 1 and em["shape"].tolist() == [0] and abs(em["area"][0] - 0.16) < 1e-6
    with pytest.raises(trb.TrbError):
        trb.HostScene.load_obj(tmp_path / "nope.obj")
    # an emissive face inside a shape whose first face is not emissive walks off the reference's emitter list
    # (integrator.cpp:53-58): rejected instead of reproduced
    (tmp_path / "bad.obj").write_text(OBJ.replace("usemtl grey", "usemtl light"))
    with pytest.raises(trb.TrbError):
        trb.HostScene.load_obj(tmp_path / "bad.obj").emitters()


def test_unsupported_material_is_rejected(trb, tmp_path):
    (tmp_path / "s.obj").write_text(OBJ)
    (tmp_path / "m.mtl").write_text(MTL.replace("illum 8", "illum 5"))
    s = trb.HostScene.load_obj(tmp_path / "s.obj")
    with pytest.raises(trb.TrbError):  # the reference builds no BSDF for illum 5 (renderer.cpp:259-270)
        s.emitters()


def test_obj_export_round_trip(trb, tmp_path):
    import scene_to_obj
    s = trb.HostScene.load_npz(GOLD / "scene_caustic.npz")
    back = trb.HostScene.load_obj(scene_to_obj.export_verified(s, tmp_path / "c.obj"))
    assert np.array_equal(back.positions.view(np.uint32), s.positions.view(np.uint32)) and np.array_equal(back.tri_v, s.tri_v)
    assert np.array_equal(back.bvh_order(), s.bvh_order())


def test_exr_writer(trb, tmp_path):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    img = (rng.random((37, 53, 3)) * 4).astype(np.float32)
    img[0, 0] = [0.0, 1.0, 65504.0]
    trb.save_exr(tmp_path / "a.exr", img)
    raw = (tmp_path / "a.exr").read_bytes()
    assert raw[:4] == bytes([0x76, 0x2F, 0x31, 0x01]) and len(raw) > 37 * 53 * 6
    back = cv2.imread(str(tmp_path / "a.exr"), cv2.IMREAD_UNCHANGED)  # BGR
    assert back is not None and back.shape == (37, 53, 3)
    assert np.allclose(back[..., ::-1], img, rtol=1e-3, atol=1e-4)  # half precision, channels B,G,R, row 0 = top


def test_parallel_bvh_build_is_the_sequential_build(trb, monkeypatch):
    """the multi-threaded host build (subtrees on their own threads, stitched in pre-order) yields the primitive order of the
    single-threaded, reference-order build (bvh.h:147-247) on a scene large enough to be split up"""
    import gen_tess_scene
    scene = gen_tess_scene.make_scene(300_000, seed=3)
    monkeypatch.setenv("TRB_BUILD_THREADS", "1")
    seq = scene.bvh_order()
    monkeypatch.setenv("TRB_BUILD_THREADS", "5")
    par = scene.bvh_order()
    assert np.array_equal(seq, par) and sorted(seq.tolist()) == list(range(scene.n_tris))


def test_checkpoint_round_trip_and_png(trb, tmp_path):
    """the checkpoint file (running-sum frame buffer + samples done + render description) round-trips, a damaged
    file is refused, `matches` tells the same render from another one, and the quick-look PNG decodes to the sRGB curve"""
    rng = np.random.default_rng(5)
    info = trb.CheckpointInfo()
    info.desc = trb.render_desc(width=40, height=30, spp=64, rr_depth=8)
    info.integrator, info.spp_done, info.scene_hash = 0, 24, 1234567
    rgb = rng.random((30, 40, 3)).astype(np.float32)
    f = tmp_path / "a.ckpt"
    trb.save_checkpoint(f, info, rgb)
    got, img = trb.load_checkpoint(f)
    assert (got.spp_done, got.desc.spp, got.desc.width, got.scene_hash) == (24, 64, 40, 1234567) and np.array_equal(img, rgb)
    assert trb.checkpoint_matches(info, got)
    other = trb.CheckpointInfo.from_buffer_copy(bytes(info))
    other.desc.seed = 7
    assert not trb.checkpoint_matches(info, other)
    other = trb.CheckpointInfo.from_buffer_copy(bytes(info))
    other.desc.spp_begin, other.desc.spp_end, other.desc.wave_samples, other.spp_done = 24, 64, 4096, 0   # what a resumed run passes
    assert trb.checkpoint_matches(info, other)
    raw = bytearray(f.read_bytes())
    raw[len(raw) // 2] ^= 0x40
    (tmp_path / "bad.ckpt").write_bytes(bytes(raw))
    with pytest.raises(trb.TrbError):
        trb.load_checkpoint(tmp_path / "bad.ckpt")
    (tmp_path / "short.ckpt").write_bytes(bytes(raw[: len(raw) // 3]))
    with pytest.raises(trb.TrbError):
        trb.load_checkpoint(tmp_path / "short.ckpt")
    with pytest.raises(trb.TrbError):
        trb.load_checkpoint(tmp_path / "missing.ckpt")
    cv2 = pytest.importorskip("cv2")
    ramp = np.linspace(0, 1.2, 40, dtype=np.float32)[None, :, None].repeat(30, 0).repeat(3, 2)
    ramp[:, :, 1] *= 0.5
    trb.save_png(tmp_path / "a.png", ramp, exposure=1.0)
    bgr = cv2.imread(str(tmp_path / "a.png"), cv2.IMREAD_UNCHANGED)
    assert bgr is not None and bgr.shape == (30, 40, 3) and bgr.dtype == np.uint8
    v = np.clip(ramp, 0, 1)
    want = np.where(v <= 0.0031308, 12.92 * v, 1.055 * np.power(v, 1 / 2.4) - 0.055)
    assert np.abs(bgr[:, :, ::-1].astype(np.float64) - want * 255).max() <= 1.0


def test_reference_loader_parity_if_available(trb):
    """Where the reference itself is present (this container), the loader is compared array by array."""
    import refh
    ref_root = "/root/reference/data/a5/cbox/mesh/cbox_mirror.obj"
    if not (refh.available() and os.path.exists(ref_root)):
        pytest.skip("reference not present on this box")
    hs = trb.HostScene.load_obj(ref_root)
    rh = refh.RefHarness.open_scene(ref_root, width=32, height=24)
    d = rh.scene_dump()
    for k in ("positions", "normals", "tri_v", "tri_n", "tri_mat", "tri_shape"):
        assert np.array_equal(getattr(hs, k), d[k]), k
    assert np.array_equal(hs.bvh_order(), rh.bvh_order())
    rh.close()


def test_cpp_host_api_compiles_and_fails_loudly_without_gpu(trb, tmp_path):
    """include/tinyrender/tinyrender.h is self-contained C++17, the CLI keeps the reference's syntax (main.cpp:160-171)
    and - on a box without a CUDA device - stops at Renderer::init instead of falling back to any CPU path."""
    import subprocess
    import torch
    root = Path(__file__).resolve().parent.parent
    pkg = root / "bidirectional-path-tracing_b200"
    src = tmp_path / "t.cpp"
    src.write_text('#include "tinyrender/tinyrender.h"\n#include <cstdio>\nint main(int c, char** v) { TinyRender::Config cfg;\n'
                   ' try { TinyRender::loadTOML(cfg, v[1]); } catch (const std::runtime_error& e) { std::printf("EXC %s\\n", e.what()); return 7; }\n'
                   ' std::printf("%d %d %d %d %s\\n", cfg.width, cfg.height, cfg.spp, cfg.integratorSettings.bdpt.rrDepth, cfg.objFile.c_str()); return 0; }\n')
    sys.path.insert(0, str(pkg))
    import build as trb_build
    exe = tmp_path / "t"
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Werror", "-I", str(root / "include"), str(src), "-o", str(exe), "-L", str(pkg), "-l:libtrb.so",
                    f"-Wl,-rpath,{pkg}", *trb_build._stdcxx_link_args(), "-lm"], check=True)
    toml = tmp_path / "s.toml"
    body = ('[input]\nobjfile = "m/x.obj"\n[camera]\neye = [ 0.0, 0.8, 3.8 ]\nat = [ 0.0, 0.8, 0.0 ]\nup = [ 0.0, 1.0, 0.0 ]\nfov = 30.0\n'
            '[film]\nwidth = 80\nheight = 60\n[renderer]\nrealtime = false\ntype = "%s"\nrrDepth = 7\nspp = 3\n')
    toml.write_text(body % "bdpt")
    r = subprocess.run([str(exe), str(toml)], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.split()[:4] == ["80", "60", "3", "7"] and r.stdout.strip().endswith("m/x.obj")
    toml.write_text(body % "nonsense")
    r = subprocess.run([str(exe), str(toml)], capture_output=True, text=True)
    assert r.returncode == 7 and "Invalid integrator type" in r.stdout  # main.cpp:109
    # the program that drives the API like the reference's own code - including a BSDF subclass built with the reference's constructor
    # signature BSDF(const WorldData&, const Config&, size_t) and the Scene / Integrator getters - compiles warning-free (it RUNS in
    # tests/test_gpu_image.py: Scene::load needs a device)
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Werror", "-I", str(root / "include"), str(root / "tests/cpp/api_check.cpp"), "-o", str(tmp_path / "api_check"),
                    "-L", str(pkg), "-l:libtrb.so", f"-Wl,-rpath,{pkg}", *trb_build._stdcxx_link_args(), "-lm"], check=True)
    cli = pkg / "tinyrender_b200"
    assert cli.exists()
    r = subprocess.run([str(cli)], capture_output=True, text=True)
    assert r.returncode != 0 and "Syntax" in r.stderr
    if not torch.cuda.is_available():
        toml.write_text(body % "bdpt")
        r = subprocess.run([str(cli), str(toml), "nogui"], capture_output=True, text=True)
        assert r.returncode != 0  # missing OBJ / no device: never a silent CPU render


@pytest.mark.parametrize("name", ["spherelight", "caustic", "tess"])
def test_shape_spheres_match_reference(trb, name):
    """Scene::getShapeCenter / getShapeRadius (renderer.cpp:294-304, 349-358) as the host builder derives them for the device
    (DirectIntegrator's sphere emitters) against the reference's own values (tools/gen_golden_direct.py), bit for bit"""
    from conftest import GOLD
    g = np.load(GOLD / f"dipaths_{name}_mis.npz")
    hs = trb.HostScene.load_npz(GOLD / f"scene_{name}.npz")
    assert np.array_equal(hs.shape_spheres().view(np.uint32), g["spheres"].view(np.uint32))


def test_spherelight_scene_is_reproducible_and_pinned(trb):
    """the sphere-light scene is generated (tools/gen_tess_scene.make_sphere_light_scene); the committed fixture is that output, and
    the host builder's BVH order and emitters on it are the reference's"""
    from conftest import GOLD
    import gen_tess_scene
    sc = gen_tess_scene.make_sphere_light_scene()
    hs = trb.HostScene.load_npz(GOLD / "scene_spherelight.npz")
    for k in ("positions", "normals", "tri_v", "tri_n", "tri_mat", "tri_shape"):
        assert np.array_equal(getattr(sc, k), getattr(hs, k)), k
    g = np.load(GOLD / "host_spherelight.npz")
    assert np.array_equal(hs.bvh_order(), g["order"])
    em = hs.emitters()
    assert em["n"] == 2 and np.array_equal(em["shape"], g["em_shape"]) and np.array_equal(em["area"].view(np.uint32), g["em_area"].view(np.uint32))
    assert np.array_equal(em["cdf"][: len(g["em_cdf"])].view(np.uint32), g["em_cdf"].view(np.uint32))


def test_textured_obj_loader(trb, tmp_path):
    """vt / map_Kd / map_Ks / PPM through the product's loader: the generated textured scene written out as OBJ + MTL + P6 files
    reads back bit-identically (Tex::load's gamma expansion and vertical flip, core.h:430-497; texture names resolved against the
    OBJ's directory, core.h:528-543); a missing image becomes the reference's 1x1 pink texel; texture options are skipped"""
    from conftest import GOLD
    import scene_to_obj
    sc = trb.HostScene.load_npz(GOLD / "scene_textured.npz")
    obj = scene_to_obj.export_verified(sc, tmp_path / "t.obj")   # raises unless every array, material and texel round-trips
    back = trb.HostScene.load_obj(obj)
    assert len(back.textures) == 4 and back.textures[0].shape == (48, 64, 3) and back.tri_t.min() == -1
    # row 0 of the decoded image is the LAST row of the file
    raw = (tmp_path / "t_tex0.ppm").read_bytes()
    body = np.frombuffer(raw[raw.index(b"255\n") + 4:], np.uint8).reshape(48, 64, 3)
    assert np.array_equal(scene_to_obj.ppm_lut()[body[::-1]].view(np.uint32), back.textures[0].view(np.uint32))
    mtl = (tmp_path / "t.mtl").read_text().replace("map_Kd t_tex0.ppm", "map_Kd -bm 0.5 -o 1 2 missing.png")
    (tmp_path / "t.mtl").write_text(mtl)
    pink = trb.HostScene.load_obj(obj)
    i = [k for k, t in enumerate(pink.textures) if t.shape == (1, 1, 3)]
    assert len(i) == 1 and np.array_equal(pink.textures[i[0]].ravel(), [1.0, 0.0, 1.0])


def test_scene_hash_follows_content(trb):
    """trb_host_scene_hash (the checkpoint's scene fingerprint) covers the CONTENT of the scene arrays: a moved vertex, a changed Kd /
    Ke / illum or a swapped face changes it even though every count stays the same (the old fingerprint held three counts)"""
    L = trb.lib()
    L.trb_host_scene_hash.restype = ctypes.c_uint64
    sc = trb.HostScene.load_npz(GOLD / "scene_caustic.npz")

    def h(scene):
        d = scene.desc()
        return int(L.trb_host_scene_hash(ctypes.byref(d)))
    base = h(sc)
    assert base != 0 and h(trb.HostScene.load_npz(GOLD / "scene_caustic.npz")) == base
    seen = {base}
    for edit in ("vertex", "kd", "ke", "illum", "face", "normal"):
        e = trb.HostScene.load_npz(GOLD / "scene_caustic.npz")
        if edit == "vertex":
            e.positions[17, 1] += np.float32(1e-3)
        elif edit == "kd":
            e.materials["kd"][0, 0] = np.float32(e.materials["kd"][0, 0] * 0.5 + 0.1)
        elif edit == "ke":
            k = int(np.flatnonzero(e.materials["ke"].sum(1) > 0)[0]); e.materials["ke"][k] *= np.float32(2)
        elif edit == "illum":
            e.materials["illum"][0] = 3 if e.materials["illum"][0] != 3 else 7
        elif edit == "face":
            e.tri_v[[5, 6]] = e.tri_v[[6, 5]]
        else:
            e.normals[3] = -e.normals[3]
        v = h(e)
        assert v not in seen, edit
        seen.add(v)
    assert h(trb.HostScene.load_npz(GOLD / "scene_hardlight.npz")) not in seen


def test_phong_lobe_cosine_is_symmetric_bit_for_bit():
    """bsdf.cuh: glossy_lobe - the value, the pdf and the reverse pdf of a glossy connection share one powf because
    dot(wi, reflect_z(wo)) and dot(wo, reflect_z(wi)) are the same float: the products commute, negation is exact, and the sums are
    taken in the same order (device code is built -fmad=false, so nothing is contracted).  Checked here in float32 on 10^6 random pairs,
    including denormals and huge values."""
    rng = np.random.default_rng(11)
    n = 1_000_000
    scale = np.float32(10.0) ** rng.integers(-30, 30, (n, 1)).astype(np.float32)
    wo = (rng.standard_normal((n, 3)).astype(np.float32) * scale).astype(np.float32)
    wi = (rng.standard_normal((n, 3)).astype(np.float32) * scale[::-1]).astype(np.float32)

    def dot_reflected(a, b):   # dot(a, reflect_z(b)) = (a.x * -b.x + a.y * -b.y) + a.z * b.z, float32 at every step
        return (a[:, 0] * (-b[:, 0]) + a[:, 1] * (-b[:, 1])) + a[:, 2] * b[:, 2]

    with np.errstate(over="ignore", invalid="ignore"):
        c1, c2 = dot_reflected(wi, wo), dot_reflected(wo, wi)
    assert c1.dtype == np.float32 and np.array_equal(c1.view(np.uint32), c2.view(np.uint32))
