"""GPU tests, level 3 and the host surface: converged images against the reference's own shipped renders, the
size-independent properties of the render call (sample-range additivity, wave-size independence), the CLI and the
C++ host API (include/tinyrender/tinyrender.h) driven like the reference's own code."""
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent
GOLD = ROOT / "tests" / "golden"
PKG = ROOT / "bidirectional-path-tracing_b200"


def _box(img, k):
    h, w, c = img.shape
    return img.reshape(h // k, k, w // k, k, c).astype(np.float64).mean((1, 3))


# tolerance per pin: box-filtered relMSE.  The shipped hard-light image (cbox_bdpt_final.exr) predates the final code:
# the unmodified reference at HEAD differs from it by relMSE 5.9e-3 (tools/gen_image_golden.py), so that pin gets 1e-2
# and the tight hard-light pin is a 96 spp render of the reference itself (image_hardlight_rr8_head.npz).
PINS = {"caustic_rr5": 2e-3, "caustic_rr8": 2e-3, "hardlight_rr8_head": 2e-3, "hardlight_rr16_head": 2e-3, "hardlight_rr8": 1e-2}


@pytest.mark.parametrize("pin", list(PINS))
def test_image_matches_reference_golden(trb, gpu_ctx, pin):
    """relMSE of a 128 spp render against the reference's converged image (box-filtered 8x8 on both sides)."""
    g = np.load(GOLD / f"image_{pin}.npz")
    ctx = gpu_ctx(str(g["scene"]))
    desc = trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=128, rr_depth=int(g["rr_depth"]))
    img, st = ctx.render_bdpt(desc)
    a, b = _box(img, int(g["box"])), g["image"].astype(np.float64)
    rel_mse = float(np.mean((a - b) ** 2 / (b ** 2 + 1e-3)))
    mean_ratio = float(img.mean() / float(g["mean"]))
    print(f"{pin}: relMSE {rel_mse:.2e}, mean ratio {mean_ratio:.4f}, {st['samples'] / st['ms_total'] / 1e3:.1f} Msamples/s")
    assert rel_mse <= PINS[pin], rel_mse
    assert abs(mean_ratio - 1) <= 0.01, mean_ratio


def test_russian_roulette_image(trb, gpu_ctx):
    """TRB_FLAG_RUSSIAN_ROULETTE at image level: (1) against a render by the reference built with its own Russian-
    roulette code (NO_RR 0; tools/gen_image_golden.py --rr) when that pin is present; (2) unbiasedness against the
    hard-capped estimator run to the same depth: roulette only re-weights the tail, so the two converge to the same
    image (cap 32 on both sides)."""
    ctx = gpu_ctx("caustic")
    kw = dict(width=800, height=600, spp=128, rr_depth=5)
    rr, st = ctx.render_bdpt(trb.render_desc(flags=trb.TRB_FLAG_RUSSIAN_ROULETTE, **kw))
    pin = GOLD / "image_caustic_rr5_roulette.npz"
    if pin.exists():
        g = np.load(pin)
        a, b = _box(rr, int(g["box"])), g["image"].astype(np.float64)
        rel_mse = float(np.mean((a - b) ** 2 / (b ** 2 + 1e-3)))
        ratio = float(rr.mean() / float(g["mean"]))
        print(f"roulette vs NO_RR-0 reference: relMSE {rel_mse:.2e}, mean ratio {ratio:.4f}")
        assert rel_mse <= 3e-3 and abs(ratio - 1) <= 0.01
    deep, st2 = ctx.render_bdpt(trb.render_desc(width=800, height=600, spp=128, rr_depth=trb.TRB_RR_DEFAULT_MAX_DEPTH))
    a, b = _box(rr, 8), _box(deep, 8)
    rel_mse = float(np.mean((a - b) ** 2 / (b ** 2 + 1e-3)))
    print(f"roulette vs hard cap {trb.TRB_RR_DEFAULT_MAX_DEPTH}: relMSE {rel_mse:.2e}, mean ratio {rr.mean() / deep.mean():.4f}; rays/sample "
          f"{(st['rays_closest'] + st['rays_any']) / st['samples']:.1f} vs {(st2['rays_closest'] + st2['rays_any']) / st2['samples']:.1f}")
    assert rel_mse <= 3e-3 and abs(rr.mean() / deep.mean() - 1) <= 0.01
    # and it is not the capped render at rr_depth 5 (which loses the deeper bounces)
    shallow, _ = ctx.render_bdpt(trb.render_desc(**kw))
    assert rr.mean() > shallow.mean() * 1.01


def test_path_tracer_image(trb, gpu_ctx):
    """the path tracer at image level: (1) against the reference's shipped cbox_bdpt_path.exr (PathTracerIntegrator, explicit,
    Russian roulette 0.95 past depth 5, 512 spp on the hard-light scene - a noisy image of a hard scene, hence the wider bound);
    (2) against BDPT on the same scene, the cross-check the reference's author made.  The two are NOT estimators of the same
    integral: recursiveExplicit drops every indirect direction that lands on the emitter (path.h:108-117), so light reaching a
    surface through a specular bounce is missing; the reference's own shipped pair differs by the same amount
    (cbox_bdpt_path.exr mean 0.1162 vs cbox_bdpt_final.exr 0.1255, ratio 0.926).
    The shipped image is itself a 512 spp render, so the statistic is (its noise) + (ours).  At 512 spp on both sides it sat AT the
    bound and moved with the random streams (1.98e-2 with the 32-bit stream keys of round 1, 2.12e-2 with the 64-bit keys); the
    device side is rendered at 2048 spp so that the bound tests the estimator, not one noise realisation."""
    ctx = gpu_ctx("hardlight")
    g = np.load(GOLD / "image_hardlight_path.npz")
    desc = trb.render_desc(width=800, height=600, spp=2048, rr_depth=int(g["rr_depth"]))
    pt, st = ctx.render_path(desc, trb.path_desc())
    a, b = _box(pt, int(g["box"])), g["image"].astype(np.float64)
    rel_mse = float(np.mean((a - b) ** 2 / (b ** 2 + 1e-3)))
    ratio = float(pt.mean() / float(g["mean"]))
    print(f"path tracer vs cbox_bdpt_path.exr: relMSE {rel_mse:.2e}, mean ratio {ratio:.4f}, {st['samples'] / st['ms_total'] / 1e3:.1f} Msamples/s, "
          f"{st['rays_closest'] / st['samples']:.1f} rays/sample")
    assert rel_mse <= 2e-2 and abs(ratio - 1) <= 0.03
    bd, _ = ctx.render_bdpt(trb.render_desc(width=800, height=600, spp=128, rr_depth=int(g["rr_depth"]), flags=trb.TRB_FLAG_RUSSIAN_ROULETTE))
    c = _box(bd, 8)
    rel2 = float(np.mean((a - c) ** 2 / (c ** 2 + 1e-3)))
    print(f"path tracer vs BDPT (roulette): relMSE {rel2:.2e}, mean ratio {pt.mean() / bd.mean():.4f}")
    assert 0.89 <= pt.mean() / bd.mean() <= 0.97


def test_sample_range_additivity(trb, gpu_ctx, scene_name):
    """disjoint sample ranges are plain summands (what sharding by sample index relies on)"""
    ctx = gpu_ctx(scene_name)
    kw = dict(width=200, height=150, spp=8, rr_depth=8)
    full, _ = ctx.render_bdpt(trb.render_desc(**kw))
    ctx.render_bdpt(trb.render_desc(spp_begin=0, spp_end=3, **kw), download=False)
    parts, _ = ctx.render_bdpt(trb.render_desc(spp_begin=3, spp_end=8, **kw), accumulate=True)
    # float atomics commute but do not associate: allow a few ulp of the largest summand per pixel
    assert np.allclose(full, parts, rtol=2e-5, atol=2e-6), float(np.abs(full - parts).max())
    assert full.mean() > 0.01


def test_render_into_the_callers_frame_buffer(trb, gpu_ctx):
    """trb_render_bdpt writes the frame into the caller's own array (the reference's RenderBuffer is caller-owned): a pinned one here,
    as bench.py's end-to-end loop uses; a wrong shape or dtype is refused before anything is launched"""
    import torch
    ctx = gpu_ctx("caustic")
    kw = dict(width=200, height=150, spp=4, rr_depth=8)
    ref, _ = ctx.render_bdpt(trb.render_desc(**kw))
    frame = torch.empty((150, 200, 3), dtype=torch.float32, pin_memory=True).numpy()
    frame[:] = -1.0
    got, st = ctx.render_bdpt(trb.render_desc(**kw), out=frame)
    assert got is frame and st["samples"] == 200 * 150 * 4
    assert np.allclose(frame, ref, rtol=2e-5, atol=2e-6) and frame.min() >= 0.0
    with pytest.raises(ValueError):
        ctx.render_bdpt(trb.render_desc(**kw), out=np.zeros((150, 200, 4), np.float32))
    with pytest.raises(ValueError):
        ctx.render_bdpt(trb.render_desc(**kw), out=np.zeros((150, 200, 3), np.float64))


def test_4k_film_sharded_by_sample_index(trb, gpu_ctx):
    """configs[3] at its film size (3840x2160, 99.5 MB frame buffer): two sample shards accumulated == one render, and the
    downsampled image agrees with an 800x600 render of the same camera (same estimator, same aspect handling)"""
    ctx = gpu_ctx("caustic")
    kw = dict(width=3840, height=2160, spp=2, rr_depth=8)
    full, st = ctx.render_bdpt(trb.render_desc(**kw))
    assert st["samples"] == 3840 * 2160 * 2
    ctx.render_bdpt(trb.render_desc(spp_begin=0, spp_end=1, **kw), download=False)
    parts, _ = ctx.render_bdpt(trb.render_desc(spp_begin=1, spp_end=2, **kw), accumulate=True)
    assert np.allclose(full, parts, rtol=2e-5, atol=2e-6), float(np.abs(full - parts).max())
    small, _ = ctx.render_bdpt(trb.render_desc(width=1920, height=1080, spp=8, rr_depth=8))
    a, b = _box(full, 48), _box(small, 24)   # both 80 x 45
    rel_mse = float(np.mean((a - b) ** 2 / (b ** 2 + 1e-3)))
    print(f"4K vs 1080p: relMSE {rel_mse:.2e}, mean ratio {full.mean() / small.mean():.4f}")
    assert rel_mse <= 3e-3 and abs(full.mean() / small.mean() - 1) <= 0.01


def test_in_library_multi_gpu_render(trb, gpu_ctx):
    """trb_render_bdpt_multi / trb_render_path_multi (needs >= 2 GPUs): sample-index shards on two devices + one ncclReduce inside the
    library == the single-device render of the same sample indices (same RNG streams; float atomics may reorder)"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    scene = trb.HostScene.load_npz(GOLD / "scene_caustic.npz")
    ctxs = [gpu_ctx("caustic"), trb.Context(1)]
    ctxs[1].upload(scene)
    desc = trb.render_desc(width=320, height=240, spp=8, rr_depth=8)
    one, _ = ctxs[0].render_bdpt(desc)
    two, st = trb.render_bdpt_multi(ctxs, desc)
    assert np.allclose(one, two, rtol=2e-5, atol=2e-6), float(np.abs(one - two).max())
    assert st["samples"] == 320 * 240 * 4          # device 0 rendered its half
    pd = trb.path_desc()
    one, _ = ctxs[0].render_path(trb.render_desc(width=320, height=240, spp=8, rr_depth=5), pd)
    two, _ = trb.render_path_multi(ctxs, trb.render_desc(width=320, height=240, spp=8, rr_depth=5), pd)
    assert np.allclose(one, two, rtol=2e-5, atol=2e-6), float(np.abs(one - two).max())
    dd = trb.direct_desc("mis", 2, 2)
    one, _ = ctxs[0].render_direct(trb.render_desc(width=320, height=240, spp=8), dd)
    two, _ = trb.render_direct_multi(ctxs, trb.render_desc(width=320, height=240, spp=8), dd)
    assert np.allclose(one, two, rtol=2e-5, atol=2e-6), float(np.abs(one - two).max())
    # a scene built on the second device by the device BVH build serves the same frame
    ctxs[1].upload(scene, build="device")
    two, _ = trb.render_bdpt_multi(ctxs, desc)
    one, _ = ctxs[0].render_bdpt(desc)
    assert np.allclose(one, two, rtol=2e-5, atol=2e-6)
    ctxs[1].close()
    # and through the host surface: `gpus = 2` in the TOML
    cv2 = pytest.importorskip("cv2")
    import tempfile
    tmp = Path(tempfile.mkdtemp(prefix="trb_multi_"))
    _, toml = _export(trb, "caustic", tmp, extra="gpus = 2\n", width=160, height=120, rr_depth=8, spp=4)
    r = subprocess.run([str(PKG / "tinyrender_b200"), str(toml), "nogui"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "NCCL is not available" not in r.stdout, r.stdout + r.stderr
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    bgr = cv2.imread(str(tmp / "scene.exr"), cv2.IMREAD_UNCHANGED)
    img, _ = ctxs[0].render_bdpt(trb.render_desc(width=160, height=120, spp=4, rr_depth=8))
    assert bgr is not None and np.allclose(bgr[:, :, ::-1], img, rtol=2e-3, atol=1e-4)


def test_wave_size_independence(trb, gpu_ctx, scene_name):
    ctx = gpu_ctx(scene_name)
    kw = dict(width=200, height=150, spp=4, rr_depth=8)
    a, _ = ctx.render_bdpt(trb.render_desc(**kw))
    b, _ = ctx.render_bdpt(trb.render_desc(wave_samples=7777, **kw))
    assert np.allclose(a, b, rtol=2e-5, atol=2e-6), float(np.abs(a - b).max())


def _export(trb, name, tmp_path, type_="bdpt", extra="", rr_prob=0.95, **toml):
    sys.path.insert(0, str(ROOT / "tools"))
    import scene_to_obj
    scene = trb.HostScene.load_npz(GOLD / f"scene_{name}.npz")
    obj = scene_to_obj.export_verified(scene, tmp_path / "scene.obj")
    t = tmp_path / "scene.toml"
    t.write_text('[input]\nobjfile = "scene.obj"\n\n[camera]\neye = [ 0.0, 0.8, 3.8 ]\nat = [ 0.0, 0.8, 0.0 ]\nup = [ 0.0, 1.0, 0.0 ]\nfov = 30.0\n\n'
                 f'[film]\nwidth = {toml["width"]}\nheight = {toml["height"]}\n\n[renderer]\nrealtime = false\ntype = "{type_}"\n'
                 f'rrDepth = {toml["rr_depth"]}\nrrProb = {rr_prob}\nspp = {toml["spp"]}\n{extra}')
    return obj, t


def test_cli_renders_exr(trb, gpu_ctx, tmp_path):
    """tinyrender_b200 scene.toml nogui == the reference's CLI contract: prints the render time, writes <stem>.exr"""
    cv2 = pytest.importorskip("cv2")
    _, toml = _export(trb, "caustic", tmp_path, width=160, height=120, rr_depth=8, spp=4)
    r = subprocess.run([str(PKG / "tinyrender_b200"), str(toml), "nogui"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Render took:" in r.stdout and "Saved EXR" in r.stdout
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    bgr = cv2.imread(str(tmp_path / "scene.exr"), cv2.IMREAD_UNCHANGED)
    assert bgr is not None and bgr.shape == (120, 160, 3)
    img, _ = gpu_ctx("caustic").render_bdpt(trb.render_desc(width=160, height=120, spp=4, rr_depth=8))
    assert np.allclose(bgr[:, :, ::-1], img, rtol=2e-3, atol=1e-4)  # half-float storage
    bad = subprocess.run([str(PKG / "tinyrender_b200")], capture_output=True, text=True)
    assert bad.returncode != 0 and "Syntax" in bad.stderr


def test_cli_checkpoint_resume(trb, gpu_ctx, tmp_path):
    """a run interrupted after its first checkpoint and started again resumes from the file and ends with the
    image of an uninterrupted run; the preview PNG is there."""
    cv2 = pytest.importorskip("cv2")
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    extra = 'checkpoint = "run.ckpt"\ncheckpointEvery = 3\npreview = "run.png"\n'
    _, toml = _export(trb, "caustic", tmp_path, extra=extra, width=160, height=120, rr_depth=8, spp=8)
    exe = [str(PKG / "tinyrender_b200"), str(toml), "nogui"]
    r1 = subprocess.run(exe, capture_output=True, text=True, timeout=300, env=dict(os.environ, TRB_STOP_AFTER_CHECKPOINTS="1"))
    assert r1.returncode == 0 and "Checkpoint: 3 of 8" in r1.stdout and "Stopped after 1" in r1.stdout, r1.stdout + r1.stderr
    info, part = trb.load_checkpoint(tmp_path / "run.ckpt")
    assert info.spp_done == 3 and info.desc.spp == 8 and part.mean() > 0
    r2 = subprocess.run(exe, capture_output=True, text=True, timeout=300)
    assert r2.returncode == 0 and "Resuming from" in r2.stdout and "3 of 8" in r2.stdout and "Checkpoint: 8 of 8" in r2.stdout, r2.stdout + r2.stderr
    assert "Checkpoint: 3 of 8" not in r2.stdout   # the first slice was not rendered again
    bgr = cv2.imread(str(tmp_path / "scene.exr"), cv2.IMREAD_UNCHANGED)
    img, _ = gpu_ctx("caustic").render_bdpt(trb.render_desc(width=160, height=120, spp=8, rr_depth=8))
    assert bgr is not None and np.allclose(bgr[:, :, ::-1], img, rtol=2e-3, atol=1e-4)
    png = cv2.imread(str(tmp_path / "run.png"), cv2.IMREAD_UNCHANGED)
    assert png is not None and png.shape == (120, 160, 3) and png.mean() > 10
    # a different render (other seed) does not pick the checkpoint up
    toml.write_text(toml.read_text() + "seed = 99\n")
    r3 = subprocess.run(exe, capture_output=True, text=True, timeout=300)
    assert r3.returncode == 0 and "Resuming from" not in r3.stdout


def test_cli_and_cpp_api_path_tracer(trb, gpu_ctx, gold, tmp_path):
    """`type = "path"` through the same host surface: the CLI writes the frame trb_render_path gives, and
    PathTracerIntegrator::render(ray, sampler) with Sampler(seed) per path returns the unmodified reference's Li
    (tests/golden/ptpaths_caustic_mis.npz: explicit, maxDepth 3, 2 emitter + 2 BSDF samples)."""
    cv2 = pytest.importorskip("cv2")
    sys.path.insert(0, str(PKG))
    import build as trb_build
    g = gold("ptpaths", "caustic_mis")
    extra = "isExplicit = true\nmaxDepth = 3\nemitterSamples = 2\nbsdfSamples = 2\n"
    _, toml = _export(trb, "caustic", tmp_path, type_="path", extra=extra, width=int(g["width"]), height=int(g["height"]),
                      rr_depth=int(g["rr_depth"]), spp=int(g["spp"]))
    r = subprocess.run([str(PKG / "tinyrender_b200"), str(toml), "nogui"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Max depth is: 3" in r.stdout and "Saved EXR" in r.stdout
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    bgr = cv2.imread(str(tmp_path / "scene.exr"), cv2.IMREAD_UNCHANGED)
    img, _ = gpu_ctx("caustic").render_path(trb.render_desc(width=int(g["width"]), height=int(g["height"]), spp=int(g["spp"]), rr_depth=int(g["rr_depth"])),
                                            trb.path_desc(True, 3, 0.95, 2, 2))
    assert bgr is not None and np.allclose(bgr[:, :, ::-1], img, rtol=2e-3, atol=1e-4)
    exe = tmp_path / "api_check"
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", str(ROOT / "include"), str(ROOT / "tests/cpp/api_check.cpp"), "-o", str(exe),
                    "-L", str(PKG), "-l:libtrb.so", f"-Wl,-rpath,{PKG}", *trb_build._stdcxx_link_args(), "-lm"], check=True)
    n = 300
    rec = np.zeros(n, np.dtype([("ray", "<f4", 8), ("seed", "<i4")]))
    rec["ray"], rec["seed"] = g["rays"][:n], g["seeds"][:n]
    (tmp_path / "in.bin").write_bytes(np.int32(n).tobytes() + rec.tobytes())
    r = subprocess.run([str(exe), str(toml), str(tmp_path / "in.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    out = np.fromfile(tmp_path / "out.bin", np.float32).reshape(n, 6)
    ref = g["li"][:n]
    scale = np.maximum(np.abs(ref).max(1, keepdims=True), 1e-6)
    ok = (np.abs(out[:, :3] - ref) <= 1e-5 * np.maximum(scale, np.abs(out[:, :3]))).all(1)
    assert ok.sum() >= n - 3, ok.sum()


def test_cpp_host_api_paths_match_reference(trb, gold, tmp_path):
    """Integrator::render(ray, sampler) through the C++ host API with Sampler(seed) per path, against the Li the
    UNMODIFIED reference returned for the same rays and seeds (tests/golden/paths_*.npz)."""
    sys.path.insert(0, str(PKG))
    import build as trb_build
    g = gold("paths", "caustic")
    _, toml = _export(trb, "caustic", tmp_path, width=int(g["width"]), height=int(g["height"]), rr_depth=int(g["rr_depth"]), spp=int(g["spp"]))
    exe = tmp_path / "api_check"
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", str(ROOT / "include"), str(ROOT / "tests/cpp/api_check.cpp"), "-o", str(exe),
                    "-L", str(PKG), "-l:libtrb.so", f"-Wl,-rpath,{PKG}", *trb_build._stdcxx_link_args(), "-lm"], check=True)
    n = 400
    rec = np.zeros(n, np.dtype([("ray", "<f4", 8), ("seed", "<i4")]))
    rec["ray"], rec["seed"] = g["rays"][:n], g["seeds"][:n]
    (tmp_path / "in.bin").write_bytes(np.int32(n).tobytes() + rec.tobytes())
    r = subprocess.run([str(exe), str(toml), str(tmp_path / "in.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    out = np.fromfile(tmp_path / "out.bin", np.float32).reshape(n, 6)
    ref = g["li"][:n]
    scale = np.maximum(np.abs(ref).max(1, keepdims=True), 1e-6)
    ok = (np.abs(out[:, :3] - ref) <= 1e-5 * np.maximum(scale, np.abs(out[:, :3]))).all(1)
    assert ok.sum() >= n - 3, ok.sum()
    assert out[:, 3].view(np.int32).sum() > n // 2  # most camera rays hit the box


def test_multi_gpu_entry_points_fail_cleanly(trb, gpu_ctx, tmp_path):
    """(1) a sample range that cannot give every device a sample is refused (an empty slice would reach the renderer as
    spp_end == 0 = "up to spp" and be over-counted); (2) with NCCL unloadable (TRB_NCCL_LIB names a file that is not there - the override
    is the only library tried) trb_render_bdpt_multi answers TRB_ERR_UNSUPPORTED with the loader's message instead of crashing in
    std::string(nullptr) - the status the host API turns into a host-side sum."""
    a, b = gpu_ctx("caustic"), trb.Context(0)
    try:
        b.upload(a.scene)
        with pytest.raises(trb.TrbError) as e:
            trb.render_bdpt_multi([a, b], trb.render_desc(width=32, height=32, spp=100, rr_depth=4, spp_begin=0, spp_end=1))
        assert e.value.code == trb.TRB_ERR_INVALID and "at least one sample per device" in str(e.value)
    finally:
        b.close()
    prog = tmp_path / "nccl_missing.py"
    prog.write_text(
        "import importlib, sys\n"
        f"sys.path.insert(0, {str(ROOT)!r})\n"
        "trb = importlib.import_module('bidirectional-path-tracing_b200')\n"
        f"sc = trb.HostScene.load_npz({str(GOLD / 'scene_caustic.npz')!r})\n"
        "a, b = trb.Context(0), trb.Context(0)\n"
        "a.upload(sc); b.upload(sc)\n"
        "try:\n"
        "    trb.render_bdpt_multi([a, b], trb.render_desc(width=32, height=32, spp=4, rr_depth=4))\n"
        "    print('NO ERROR')\n"
        "except trb.TrbError as e:\n"
        "    print('CODE', e.code, str(e))\n")
    env = dict(os.environ, TRB_NCCL_LIB=str(tmp_path / "no_such_libnccl.so"))
    r = subprocess.run([sys.executable, str(prog)], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert f"CODE {trb.TRB_ERR_UNSUPPORTED}" in r.stdout and "no_such_libnccl.so" in r.stdout, r.stdout


def test_patched_reference_renders_on_the_device(trb, gpu_ctx, tmp_path):
    """INTEGRATION.md section B, compiled and run (`make -C oracle ref_b200` -> oracle/_ref/tinyrender_ref_b200): the reference's OWN
    main(), loadTOML, tinyobj loader, Scene::load and saveEXR, with BDPTIntegrator::renderFrame bound to libtrb.so by the three edits
    the document lists.  The frame must come from the device ("B200 path:" line, no CPU pixel loop), equal the library's own render of
    the same description (same seed => same sample streams; half-float storage), and match the reference's shipped converged image."""
    cv2 = pytest.importorskip("cv2")
    exe = ROOT / "oracle" / "_ref" / "tinyrender_ref_b200"
    if not exe.exists():
        pytest.skip("oracle/_ref/tinyrender_ref_b200 not built (make -C oracle ref_b200, where /root/reference exists)")
    g = np.load(GOLD / "image_caustic_rr8.npz")
    W, H, rr = int(g["width"]), int(g["height"]), int(g["rr_depth"])
    _, toml = _export(trb, "caustic", tmp_path, width=W, height=H, rr_depth=rr, spp=128)
    r = subprocess.run([str(exe), str(toml), "nogui"], capture_output=True, text=True, timeout=900, env=dict(os.environ, TRB_SEED="4242"))
    assert r.returncode == 0, r.stdout + r.stderr
    assert "B200 path:" in r.stdout and "Saved EXR" in r.stdout, r.stdout + r.stderr
    assert "falling back" not in r.stdout + r.stderr
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    bgr = cv2.imread(str(tmp_path / "scene.exr"), cv2.IMREAD_UNCHANGED)
    assert bgr is not None and bgr.shape == (H, W, 3)
    ref_img = bgr[:, :, ::-1].astype(np.float32)
    mine, _ = gpu_ctx("caustic").render_bdpt(trb.render_desc(width=W, height=H, spp=128, rr_depth=rr, seed=4242))
    # same description and seed: the same sample streams (float atomics may reorder sums; the EXR stores halves)
    assert np.allclose(ref_img, mine, rtol=4e-3, atol=2e-4), float(np.abs(ref_img - mine).max())
    a, b = _box(ref_img, int(g["box"])), g["image"].astype(np.float64)
    rel_mse = float(np.mean((a - b) ** 2 / (b ** 2 + 1e-3)))
    ratio = float(ref_img.mean() / float(g["mean"]))
    print(f"patched reference on the device vs cbox_bdpt_glass_8depth.exr: relMSE {rel_mse:.2e}, mean ratio {ratio:.4f}")
    assert rel_mse <= 2e-3 and abs(ratio - 1) <= 0.01
