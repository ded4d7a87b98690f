"""tools/time_upload.py -- time-to-first-pixel of the 10 M-triangle scene: trb_scene_upload with the reference-order
BVH built on the host cores vs on the device (csrc/kernels/bvh_build.cuh), then a 1 spp frame.
    python tools/time_upload.py [tris]"""
import importlib
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tools"))
import gen_tess_scene  # noqa: E402

trb = importlib.import_module("bidirectional-path-tracing_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
t0 = time.perf_counter()
sc = gen_tess_scene.make_scene(n, seed=1)
print(f"scene: {sc.n_tris} triangles generated in {time.perf_counter() - t0:.2f} s", flush=True)
t0 = time.perf_counter(); order = sc.bvh_order(); t_host_build = time.perf_counter() - t0
print(f"host reference-order BVH build alone: {t_host_build:.2f} s", flush=True)
for mode in ("host", "device", "host", "device"):
    ctx = trb.Context(0)
    t0 = time.perf_counter(); ctx.upload(sc, build=mode); t_up = time.perf_counter() - t0
    t0 = time.perf_counter(); img, st = ctx.render_bdpt(trb.render_desc(width=1024, height=1024, spp=1, rr_depth=8)); t_frame = time.perf_counter() - t0
    t0 = time.perf_counter(); ctx.refresh_scene(); t_ref = time.perf_counter() - t0
    print(f"build on {mode:6s}: trb_scene_upload {t_up:.3f} s, first 1 spp frame {t_frame:.3f} s (device {st['ms_total']:.1f} ms), "
          f"first trb_scene_refresh {t_ref:.3f} s, {ctx.scene_bytes() / 1e9:.2f} GB device scene, mean {img.mean():.4f}", flush=True)
    ctx.close()
