"""GPU tests of SURVEY 8(f) N4, bitmap textures (BitmapTexture3f, core.h:525-585) on the device: the generated textured scene
(map_Kd on Diffuse / Phong, map_Kd + map_Ks on Mixture; wrapped and negative texture coordinates) through the three integrators of
the C ABI against the UNMODIFIED reference, which read the exported OBJ + MTL + PPM files itself (tools/gen_golden_textured.py)."""
import tempfile
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL = 1e-5
GOLD = Path(__file__).parent / "golden"


def _rays(trb, r8):
    r = np.zeros(len(r8), trb.RAY_DTYPE)
    r["o"], r["d"], r["min_t"], r["max_t"] = r8[:, 0:3], r8[:, 3:6], r8[:, 6], r8[:, 7]
    return r


def _agree(li, ref):
    scale = np.maximum(np.abs(ref).max(1, keepdims=True), 1e-6)
    return (np.abs(li - ref) <= REL * np.maximum(scale, np.abs(li))).all(1)


def test_textured_bdpt_paths_match_reference(trb, gpu_ctx, gold, parity_record):
    g = gold("paths", "textured")
    desc = trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=int(g["rr_depth"]))
    li, pix, rgb, draws = gpu_ctx("textured").bdpt_paths(desc, _rays(trb, g["rays"]), g["rng"])
    from test_gpu_parity import _paths_agree   # Li + the splat pixel -> rgb map (the reference lists splats by pixel)
    ok = _paths_agree(g["li"], g["splat_pix"], g["splat_rgb"], li, pix, rgb) & (draws == g["draws"])
    print(f"textured BDPT: {ok.sum()}/{len(ok)} paths agree; mean Li {li.mean():.4f} (reference {g['li'].mean():.4f})")
    parity_record("bdpt_paths_golden/textured", ok.sum(), len(ok))
    assert ok.sum() >= len(ok) - 4 and (li > 0).any()   # measured: 2000/2000
    # the textures matter: the same paths on the scene with its bitmaps removed differ
    plain = trb.HostScene.load_npz(GOLD / "scene_textured.npz")
    plain.materials["tex"] = 0
    ctx = trb.Context(0)
    ctx.upload(plain)
    li0, _, _, _ = ctx.bdpt_paths(desc, _rays(trb, g["rays"]), g["rng"])
    ctx.close()
    assert (~_agree(li0, g["li"])).mean() > 0.3


def test_textured_path_and_direct_match_reference(trb, gpu_ctx, gold, parity_record):
    oracle_port = pytest.importorskip("oracle_port")
    ctx = gpu_ctx("textured")
    g = gold("ptpaths", "textured_mis")
    desc = trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=int(g["rr_depth"]))
    pd = trb.path_desc(bool(g["is_explicit"]), int(g["max_depth"]), float(g["rr_prob"]), int(g["emitter_samples"]), int(g["bsdf_samples"]))
    streams = np.stack([oracle_port.sampler_stream(int(s), int(g["draws"].max()) + 16) for s in g["seeds"]])
    li, draws = ctx.path_paths(desc, pd, _rays(trb, g["rays"]), streams)
    ok = _agree(li, g["li"]) & (draws == g["draws"])
    print(f"textured path tracer: {ok.sum()}/{len(ok)} agree")
    parity_record("path_tracer/textured/mis", ok.sum(), len(ok))
    assert ok.sum() >= len(ok) - 3   # measured: 1200/1200
    g = gold("dipaths", "textured_mis")
    streams = np.stack([oracle_port.sampler_stream(int(s), int(g["draws"].max()) + 4) for s in g["seeds"]])
    li, draws = ctx.direct_paths(desc, trb.direct_desc("mis", int(g["emitter_samples"]), int(g["bsdf_samples"])), _rays(trb, g["rays"]), streams)
    ok = _agree(li, g["li"]) & (draws == g["draws"])
    print(f"textured direct integrator: {ok.sum()}/{len(ok)} agree")
    parity_record("direct_paths/textured/mis", ok.sum(), len(ok))
    assert ok.sum() >= len(ok) - 3   # measured: 1500/1500


def _box(img, k):
    h, w, _ = img.shape
    return img[: h - h % k, : w - w % k].reshape(h // k, k, w // k, k, 3).mean((1, 3)).astype(np.float64)


def test_textured_image_vs_reference_renderer(trb, gpu_ctx):
    """level 3: the device's BDPT render of the textured scene against the UNMODIFIED reference's own threaded renderer on the box's host
    cores (it loads the PPM files itself): 8x8 box-filtered relMSE and the mean"""
    refh = pytest.importorskip("refh")
    if not refh.available():
        pytest.skip("oracle/_ref/libtr_ref.so not built")
    import scene_to_obj
    scene = trb.HostScene.load_npz(GOLD / "scene_textured.npz")
    obj = scene_to_obj.export_verified(scene, Path(tempfile.mkdtemp(prefix="trb_tx_")) / "scene.obj")
    w, h, spp_ref = 200, 150, 48
    rh = refh.RefHarness.open_scene(obj, width=w, height=h, spp=spp_ref, rr_depth=6)
    ref, sec = rh.render()
    rh.close()
    img, st = gpu_ctx("textured").render_bdpt(trb.render_desc(width=w, height=h, spp=512, rr_depth=6))
    a, b = _box(img, 8), _box(ref, 8)
    rel_mse = float(np.mean((a - b) ** 2 / (b ** 2 + 1e-2)))
    ratio = float(img.mean() / ref.mean())
    print(f"textured BDPT image: relMSE {rel_mse:.2e}, mean ratio {ratio:.4f}; reference {w * h * spp_ref / sec / 1e6:.3f} Msamples/s on the host, "
          f"device {st['samples'] / st['ms_total'] / 1e3:.1f} Msamples/s")
    assert rel_mse <= 5e-3 and abs(ratio - 1) <= 0.02


def test_textured_upload_rejects_what_the_reference_cannot_run(trb):
    scene = trb.HostScene.load_npz(GOLD / "scene_textured.npz")
    ctx = trb.Context(0)
    bad = trb.HostScene.load_npz(GOLD / "scene_textured.npz")
    bad.tri_t[:] = -1   # a textured material on faces without vt: the reference indexes texcoords[-1]
    with pytest.raises(trb.TrbError):
        ctx.upload(bad)
    bad = trb.HostScene.load_npz(GOLD / "scene_textured.npz")
    bad.materials["tex"][1] = 9
    with pytest.raises(trb.TrbError):
        ctx.upload(bad)
    ctx.upload(scene)
    assert ctx.scene_bytes() > 0
    ctx.close()
