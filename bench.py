#!/usr/bin/env python
"""bench.py -- BDPT throughput on B200 (metric of BASELINE.json: Msamples/s, with Mrays/s beside it).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU renderer on the host cores

A step = one render pass of the hot path over one batch: the caustic scene (BASELINE.json configs[1]:
glass sphere + area light, 1024x1024, rrDepth 8) at --spp samples per pixel on every GPU (weak scaling by
sample index: GPU g renders sample indices [g*spp, (g+1)*spp), one NCCL reduce of the frame buffers ends the
step).  `value` is device-timed with the scene resident in HBM; `e2e` goes through the host-buffer C-ABI call
(scene H2D from pinned memory + render + frame buffer D2H inside the timed region).
One JSON line is printed by rank 0.
"""
from __future__ import annotations

import argparse
import importlib
import gc
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
for p in (ROOT, ROOT / "oracle", ROOT / "tools"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))
PKG = "bidirectional-path-tracing_b200"
GOLD = ROOT / "tests" / "golden"
WORKLOADS = {
    "caustic": dict(scene="scene_caustic.npz", width=1024, height=1024, rr_depth=8,
                    desc="caustic scene (glass sphere + area light, 32536 tris) 1024x1024 BDPT rrDepth 8"),
    "hardlight": dict(scene="scene_hardlight.npz", width=1024, height=1024, rr_depth=16,
                      desc="hard-light scene (occluded emitter, 36404 tris) 1024x1024 BDPT rrDepth 16"),
    "c1_glass": dict(scene="scene_caustic.npz", width=800, height=600, rr_depth=8,
                     desc="configs[0] as shipped: data/a5/cbox/tinyrender/cbox_bdpt_glass.toml (cbox_mirror.obj, 800x600, rrDepth 8; 256 spp = --spp 256)"),
    "c1_bonus": dict(scene="scene_hardlight.npz", width=800, height=600, rr_depth=2,
                     desc="configs[0] as shipped: data/a5/bonus_bdpt/tinyrender/cbox_bdpt.toml (bonus_bdpt/mesh/cbox.obj, 800x600, rrDepth 2; 512 spp = --spp 512)"),
    "cbox4k": dict(scene="scene_caustic.npz", width=3840, height=2160, rr_depth=8,
                   desc="configs[3]: Cornell box (cbox_mirror.obj) 3840x2160 BDPT rrDepth 8, partitioned by sample index, one reduce of the "
                        "99.5 MB frame buffer (4096 spp in total = --spp 512 on 8 GPUs)"),
    "tess": dict(scene=None, tris=10_000_000, width=1024, height=1024, rr_depth=8,
                 desc="synthetic tessellated scene (tools/gen_tess_scene.py --tris 10000000 --seed 1: three displaced spheres, mirror/glass/"
                      "phong + Mixture floor) 1024x1024 BDPT rrDepth 8"),
}
METRIC, UNIT = "bdpt_msamples_per_s", "Msamples/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="caustic", choices=list(WORKLOADS))
    ap.add_argument("--spp", type=int, default=16, help="samples per pixel per GPU per step")
    ap.add_argument("--wave", type=int, default=0, help="samples per wavefront (0 = library default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the short configs[2] / configs[4] passes of the default line")
    ap.add_argument("--no-visits", action="store_true", help="skip the counters-build pass (A/B runs: no roofline figure)")
    ap.add_argument("--_inproc", type=int, default=0, help=argparse.SUPPRESS)   # internal: N contexts in THIS process, trb_render_bdpt_multi
    ap.add_argument("--_visits", action="store_true", help=argparse.SUPPRESS)  # internal: counters build, prints visit stats
    return ap.parse_args()


# ---- clocks (recipe: /opt/skills/guides/B200_PROFILING.md) ---------------------------------------------------
class ClockSampler:
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, device_index: int):
        self.idx, self.proc, self.lines = device_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.idx)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def load_scene(trb, wl, cpu=False):
    if wl["scene"] is not None:
        return trb.HostScene.load_npz(GOLD / wl["scene"])
    import gen_tess_scene
    # the CPU reference needs one heap object per triangle and a text OBJ: it is timed on a 1 M-triangle instance
    return gen_tess_scene.make_scene(1_000_000 if cpu else wl["tris"], seed=1)


def measured_peak():
    f = ROOT / "MEASURED_PEAKS.json"
    if f.exists():
        try:
            return float(json.loads(f.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---- the reference's CPU renderer (oracle/_ref) on the host cores ----------------------------------------------
def cpu_reference_step_factory(workload: dict, budget_s: float = 8.0, verified: bool = True):
    """Returns (run_step, info): run_step() renders one bounded sample of the workload with the reference's own
    threaded renderer and returns (samples, seconds).  Film is the workload's unless a calibration run says one
    step would take much longer than budget_s, then it is halved (same scene, camera, rrDepth).
    verified=False (the --impl reference arm): the OBJ the reference reads is written from the fixture arrays by pure Python and NOT
    parsed back through the product's loader, so that libtrb.so is never mapped in that process (importing the package only declares
    ctypes structures; the library is loaded on first use)."""
    import refh
    import scene_to_obj
    trb = importlib.import_module(PKG)
    scene = load_scene(trb, workload, cpu=True)
    tmp = Path(tempfile.mkdtemp(prefix="trb_ref_"))
    obj = scene_to_obj.export_verified(scene, tmp / "scene.obj") if verified else scene_to_obj.write_obj(scene, tmp / "scene.obj", digits=9)
    hw = os.cpu_count() or 8
    threads = max(hw - 4, 1) + 1  # parallelfor.h:30-65: hw-4 workers plus the calling thread for the remainder
    if refh.available():
        cal = refh.RefHarness.open_scene(obj, width=128, height=128, spp=1, rr_depth=workload["rr_depth"])
        _, sec = cal.render()
        cal.close()
        rate = 128 * 128 / max(sec, 1e-6)
        w, h = workload["width"], workload["height"]
        while w * h / rate > 2.5 * budget_s and w > 128:
            w, h = w // 2, h // 2
        rh = refh.RefHarness.open_scene(obj, width=w, height=h, spp=1, rr_depth=workload["rr_depth"])

        def run(want_image=False):
            img, s = rh.render()
            return (w * h, s, img) if want_image else (w * h, s)
        info = dict(film=(w, h), kind="reference", cores=threads, sample=f"{w}x{h} film, 1 spp, rrDepth {workload['rr_depth']}, whole frame through "
                    f"Renderer::render of the unmodified reference (oracle/_ref), {threads} threads")
        return run, info
    import oracle_port
    w, h = 256, 256
    orc = oracle_port.Oracle(scene)

    def run(want_image=False):
        t0 = time.perf_counter()
        img, _ = orc.render(trb.render_desc(width=w, height=h, spp=1, rr_depth=workload["rr_depth"]), threads=hw)
        s = time.perf_counter() - t0
        return (w * h, s, img) if want_image else (w * h, s)
    info = dict(film=(w, h), kind="port", cores=hw, sample=f"{w}x{h} film, 1 spp, rrDepth {workload['rr_depth']}, oracle/bdpt_oracle.cpp, {hw} threads")
    return run, info


def rel_mse(img, ref):
    """per-pixel relative MSE, mean over pixels and channels of (I - R)^2 / (R^2 + 1e-2)"""
    a, b = np.asarray(img, np.float64), np.asarray(ref, np.float64)
    return float(np.mean((a - b) ** 2 / (b * b + 1e-2)))


def equal_time_relmse(trb, ctx, wl, w, h, cpu_img, cpu_s, gpu_rate):
    """BASELINE.json's "relMSE at equal time" (SURVEY 8(d)): the reference's CPU frame took cpu_s seconds; render on
    the GPU for about as long (spp chosen from the measured device rate), and compare both against a converged GPU image
    (independent seed, >= 8x the equal-time spp) that passes level-3 parity (tests/test_gpu_image.py)."""
    spp_gpu = int(max(1, min(round(cpu_s * gpu_rate / (w * h)), 4096)))
    spp_ref = int(min(max(8 * spp_gpu, 1024), 16384))
    t0 = time.perf_counter()
    img, _ = ctx.render_bdpt(trb.render_desc(width=w, height=h, spp=spp_gpu, rr_depth=wl["rr_depth"]))
    gpu_s = time.perf_counter() - t0
    ref, _ = ctx.render_bdpt(trb.render_desc(width=w, height=h, spp=spp_ref, rr_depth=wl["rr_depth"], seed=977))
    return {"film": f"{w}x{h}", "cpu_seconds": cpu_s, "cpu_spp": 1, "relmse_cpu": rel_mse(cpu_img, ref), "gpu_seconds": gpu_s,
            "gpu_spp": spp_gpu, "relmse_gpu": rel_mse(img, ref), "reference_image": f"this library, {spp_ref} spp, seed 977",
            "definition": "mean over pixels and channels of (I - R)^2 / (R^2 + 1e-2); the converged image's own noise "
                          "(1/8 of the GPU image's variance or less) is included in both figures"}


def cpu_port_all_cores(trb, wl, w, h):
    """BASELINE.md section 3, optional second row: the race-free CPU ceiling.  The reference shares one mt19937 between its
    worker threads (renderer.cpp:155) and leaves four cores idle (parallelfor.h:30-32); the restatement (oracle/bdpt_oracle.cpp,
    pinned bit-exact to the reference per path) gives every thread its own Sampler and uses every core."""
    import oracle_port
    hw = os.cpu_count() or 8
    orc = oracle_port.Oracle(load_scene(trb, wl, cpu=True))
    t0 = time.perf_counter()
    orc.render(trb.render_desc(width=w, height=h, spp=1, rr_depth=wl["rr_depth"]), threads=hw)
    sec = time.perf_counter() - t0
    orc.close()
    return {"value": w * h / sec / 1e6, "unit": UNIT, "cores": hw, "kind": "port",
            "sample": f"{w}x{h} film, 1 spp, rrDepth {wl['rr_depth']}, oracle/bdpt_oracle.cpp with one Sampler per thread, {hw} threads"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = WORKLOADS[args.workload]
    run, info = cpu_reference_step_factory(wl, verified=False)
    info.pop("film", None)
    for _ in range(args.warmup):
        run()
    tot_s, tot_n = 0.0, 0
    for _ in range(args.steps):
        n, s = run()
        tot_n += n; tot_s += s
    v = tot_n / tot_s / 1e6
    info = dict(info, value=v, unit=UNIT)
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": 1e3 * tot_s / max(args.steps, 1), "higher_is_better": True,
                      "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                      "config": {"workload": wl["desc"], "spp_per_step": 1, "timing": "reference's own wall clock around Renderer::render",
                                 "basis": "per sample; CPU arm 1 spp per step (pixel centres: no jitter draws), GPU arm --spp per step"},
                      "cpu_baseline": info, "gpu_launches": 0, "native_libraries": mapped_libraries(),
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def mapped_libraries():
    """the repo's shared objects mapped into this process (evidence for which arm ran what)"""
    out = set()
    try:
        for ln in open("/proc/self/maps"):
            f = ln.split()[-1]
            if f.endswith(".so") and str(ROOT) in f:
                out.add(str(Path(f).relative_to(ROOT)))
    except OSError:
        pass
    return sorted(out)


class quiet_stdout:
    """the reference chats on C-level stdout; keep this process's stdout a single JSON line"""
    def __enter__(self):
        sys.stdout.flush()
        self.saved = os.dup(1)
        os.dup2(2, 1)

    def __exit__(self, *a):
        sys.stdout.flush()
        os.dup2(self.saved, 1)
        os.close(self.saved)


def visits_subprocess(workload, spp, wave) -> dict | None:
    """node / triangle visit counts per ray class from the counters build (untimed, one step)."""
    cnt = ROOT / PKG / "libtrb_counters.so"
    if not cnt.exists():
        return None
    env = dict(os.environ, TRB_LIB=str(cnt))
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    try:
        r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--_visits", "--workload", workload, "--spp", str(min(spp, 4)),
                            "--wave", str(wave)], env=env, capture_output=True, text=True, timeout=600)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        sys.stderr.write(f"visit counting failed: {e}\n")
        return None


L2_BYTES = 126e6


def serial_pass(trb, torch, ctx, fb, flush, wl, W, H, total_spp, s_begin, s_end, wave):
    """one untimed-for-`value` pass with TRB_FLAG_NO_OVERLAP (every kernel alone on one stream): per-kind kernel times that are the
    kernels' own; the second of two passes is returned"""
    d = trb.render_desc(width=W, height=H, spp=total_spp, rr_depth=wl["rr_depth"], spp_begin=s_begin, spp_end=s_end, wave_samples=wave,
                        flags=trb.TRB_FLAG_NO_OVERLAP)
    ser = {}
    for _ in range(2):
        flush.zero_(); torch.cuda.synchronize()
        ser = ctx.render_bdpt_device(d, fb.data_ptr())
    return ser


def kernel_metrics(workload, kernel):
    """ncu-measured figures of a kernel on a workload from the committed summary profiles/kernel_metrics.json (tools/kernel_metrics.py
    over an ncu launch list of tools/prof_render.py): DRAM bytes and duration per launch, issue-slot utilisation, active threads per
    instruction, L1/LSU busy, L2 hit rate"""
    f = ROOT / "profiles" / "kernel_metrics.json"
    try:
        t = json.loads(f.read_text())[workload]
        return t["kernels"][kernel], f"profiles/kernel_metrics.json ({t['source']}, {t['spp']} spp, ncu --clock-control none)"
    except Exception:  # noqa: BLE001
        return None, None


def roofline_block(workload, wl, ser, vis, samples_per_pass, wave, scene_bytes):
    """achieved = ALGORITHMIC bytes of the dominant kernel / its own time.  Per ray: 32 B ray + 16 B result or pending
    contribution + 128 B per visited node of the 4-wide tree + 48 B per tested triangle (+ 64 B per binary node for the rays traced on the
    reference's own tree), visit counts from the -DTRB_COUNTERS twin library on the same workload.  The kernel is HBM-bound only when the
    acceleration structure exceeds the L2 (configs[4]); on the L2-resident Cornell scenes the line says `bound: "issue"` and carries the
    ncu issue-slot / LSU figures beside the fraction, which there measures cache re-use against the HBM peak, not DRAM traffic."""
    peak, how = measured_peak()
    any_dom = ser["ms_trace_any"] >= ser["ms_trace_closest"]
    dom = "k_trace_wide<1>" if any_dom else "k_trace_wide<0>"   # <1>: occlusion + resolve of the pending contribution, <0>: closest hit
    K = wl["rr_depth"] - 1
    waves = -(-samples_per_pass // max(int(ser.get("wave_samples", 0)) or (wave if wave > 0 else 1 << 22), 1))   # the wave size the library used
    launches = 2 * K * waves   # per wave: K light + K eye shadow batches (any) / 1 camera + K light + (K - 1) eye (closest)
    hbm_bound = scene_bytes > L2_BYTES
    roof = {"bound": "hbm" if hbm_bound else "issue", "kernel": dom, "achieved": None, "peak": peak, "unit": "GB/s", "frac": None, "traffic": None,
            "peak_source": how, "scene_bytes": int(scene_bytes)}
    nrays, ms = (ser["rays_any"], ser["ms_trace_any"]) if any_dom else (ser["rays_closest"], ser["ms_trace_closest"])
    roof.update(rays_per_launch=nrays / launches, ms_per_launch=ms / launches,
                deferred_to_binary_tree_frac=(ser.get("rays_deferred_any", 0) if any_dom else ser.get("rays_deferred", 0)) / max(nrays, 1))
    if vis:
        tag = "_any" if any_dom else ""
        nv = max(vis["rays_any" if any_dom else "rays_closest"], 1)
        nodes, tris = vis["node_visits" + tag] / nv, vis["tri_tests" + tag] / nv
        bnodes, btris = vis["node_visits_bin" + tag] / nv, vis["tri_tests_bin" + tag] / nv
        per_ray = 32 + 16 + 128 * nodes + 48 * tris + 64 * bnodes + 48 * btris
        achieved = per_ray * nrays / (ms * 1e-3) / 1e9
        roof.update(achieved=achieved, frac=achieved / peak, bytes_per_ray=per_ray, bytes_per_launch=per_ray * nrays / launches,
                    nodes_per_ray=nodes, tris_per_ray=tris, binary_nodes_per_ray=bnodes)
    km, src = kernel_metrics(workload, dom)
    if km:
        roof["traffic"] = km.get("dram_bytes_per_launch")
        roof["traffic_source"] = src
        if km.get("dram_bytes_per_launch") and km.get("us_per_launch"):
            roof["dram_frac"] = km["dram_bytes_per_launch"] / (km["us_per_launch"] * 1e-6) / 1e9 / peak   # both from the same ncu pass
        for k_out, k_in in (("issue_slot_frac", "issue_slots_busy_pct"), ("active_threads_per_inst", "threads_per_inst"), ("l1_lsu_busy_frac", "lsu_busy_pct"),
                            ("l2_hit_frac", "l2_hit_pct")):
            if km.get(k_in) is not None:
                roof[k_out] = km[k_in] / 100.0 if k_in.endswith("_pct") else km[k_in]
    roof["note"] = ("algorithmic bytes = rays x (32 B ray + 16 B result/pending + 128 B x wide nodes visited + 48 B x triangles tested); the metric counts "
                    "every visit at full size, so a tree that visits MORE nodes scores higher: Msamples/s is the quantity optimised, this one is reported. " +
                    ("acceleration structure (%.2f GB) exceeds the 126 MB L2: the HBM-bound case" % (scene_bytes / 1e9) if hbm_bound else
                     "this scene (%.1f MB) is L1/L2-resident: the kernel is bound by instruction issue and the L1/LSU pipe (ncu figures beside), DRAM moves a few %% "
                     "of the algorithmic bytes, and a fraction near or above 1 is cache re-use, not HBM efficiency" % (scene_bytes / 1e6)))
    return roof


def extra_block(trb, torch, name, spp, fb, flush, args):
    """a short single-GPU pass of another configuration (configs[2] hard light, configs[4] tessellated 10 M) with its own scene upload:
    device-timed value, end-to-end value through the host-buffer call, and the roofline block of its dominant kernel"""
    wl = WORKLOADS[name]
    W, H = wl["width"], wl["height"]
    ctx = trb.Context(torch.cuda.current_device())
    try:
        scene = load_scene(trb, wl)
        t0 = time.perf_counter()
        ctx.upload(scene)
        up = time.perf_counter() - t0
        desc = trb.render_desc(width=W, height=H, spp=spp, rr_depth=wl["rr_depth"], wave_samples=args.wave)
        steps, warm = 3, 3
        for _ in range(warm):
            ctx.render_bdpt_device(desc, fb.data_ptr())
        tot, agg = 0.0, {}
        for _ in range(steps):
            flush.zero_(); torch.cuda.synchronize()
            st = ctx.render_bdpt_device(desc, fb.data_ptr())
            tot += st["ms_total"]
            for k, v in st.items():
                agg[k] = agg.get(k, 0) + v
        samples = W * H * spp
        blk = {"workload": wl["desc"], "value": samples * steps / (tot * 1e-3) / 1e6, "unit": UNIT, "spp_per_step": spp, "steps": steps, "warmup": warm,
               "ms_per_step": tot / steps, "rays_per_sample": (agg["rays_closest"] + agg["rays_any"]) / max(agg["samples"], 1),
               "scene_upload": {"seconds": round(up, 3), "triangles": int(scene.n_tris), "built_on": "device" if ctx.built_on_device() else "host"}}
        # end to end: the scene goes up ONCE (what a caller does), then render + frame buffer D2H per step
        frame = torch.empty((H, W, 3), dtype=torch.float32, pin_memory=True).numpy()   # the caller's frame buffer, pinned
        ctx.render_bdpt(desc, out=frame)   # untimed: first use of the buffer
        torch.cuda.synchronize()
        gc.collect(); gc.disable()   # (a generation-2 collection inside a three-step wall-clock loop shows as a 100 ms step)
        t0 = time.perf_counter()
        step_ms = []
        for _ in range(steps):
            ts = time.perf_counter()
            ctx.render_bdpt(desc, out=frame)
            step_ms.append(round((time.perf_counter() - ts) * 1e3, 2))
        e2e_s = time.perf_counter() - t0
        gc.enable()
        blk["e2e"] = {"value": samples * steps / e2e_s / 1e6, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": W * H * 3 * 4, "step_ms": step_ms,
                      "timing": "host wall clock around trb_render_bdpt (frame buffer out into the caller's pinned host array); the scene (%.2f GB) was uploaded once before" % (ctx.scene_bytes() / 1e9)}
        ser = serial_pass(trb, torch, ctx, fb, flush, wl, W, H, spp, 0, spp, args.wave)
        blk["stage_ms_serial"] = {k: ser[k] for k in ("ms_total", "ms_trace_closest", "ms_trace_any", "ms_logic")}
        vis = None if args.no_visits else visits_subprocess(name, spp, args.wave)
        blk["roofline"] = roofline_block(name, wl, ser, vis, samples, args.wave, ctx.scene_bytes())
        return blk
    finally:
        ctx.close()


def run_inproc_child(args):
    """one process, N contexts (one per device), trb_render_bdpt_multi: sample-index shards + ONE ncclReduce inside the library - the path
    the C++ host API and the CLI use for `gpus = N`.  Wall clock around the call, frame buffer download included."""
    trb = importlib.import_module(PKG)
    n = args._inproc
    wl = WORKLOADS[args.workload]
    scene = load_scene(trb, wl)
    ctxs = [trb.Context(g) for g in range(n)]
    for c in ctxs:
        c.upload(scene)
    W, H = wl["width"], wl["height"]
    desc = trb.render_desc(width=W, height=H, spp=args.spp * n, rr_depth=wl["rr_depth"], wave_samples=args.wave)
    for _ in range(max(args.warmup, 1)):
        trb.render_bdpt_multi(ctxs, desc)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        img, _ = trb.render_bdpt_multi(ctxs, desc)
    sec = time.perf_counter() - t0
    print(json.dumps({"value": W * H * args.spp * n * args.steps / sec / 1e6, "unit": UNIT, "n_gpus": n, "steps": args.steps, "ms_per_step": 1e3 * sec / args.steps,
                      "image_mean": float(img.mean()), "d2h_bytes_per_step": W * H * 3 * 4,
                      "timing": "host wall clock around trb_render_bdpt_multi (one process, one context per device, one ncclReduce, host frame buffer out)"}))


def inproc_subprocess(args, n) -> dict:
    """the in-library multi-GPU path, measured by rank 0 in a subprocess of its own with a hard timeout (the other ranks wait at a barrier)"""
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "TORCHELASTIC_RUN_ID",
                                                            "GROUP_RANK", "ROLE_RANK", "LOCAL_WORLD_SIZE", "GROUP_WORLD_SIZE", "ROLE_WORLD_SIZE")}
    try:
        r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--_inproc", str(n), "--workload", args.workload, "--spp", str(args.spp), "--wave", str(args.wave),
                            "--steps", str(min(args.steps, 3)), "--warmup", "2"], env=env, capture_output=True, text=True, timeout=240)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)[:300]}


def run_ours(args):
    import torch
    import torch.distributed as dist
    trb = importlib.import_module(PKG)
    par = importlib.import_module(PKG + ".parallel")
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[args.workload]
    torch.cuda.set_device(local)
    cpu_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        if not args.no_extras:
            cpu_group = dist.new_group(backend="gloo")
    scene = load_scene(trb, wl)
    ctx = trb.Context(local)
    t_up = time.perf_counter()
    ctx.upload(scene)   # trb_scene_upload: BVH::build on the device from 65536 triangles on, on the host cores below
    upload_info = {"seconds": round(time.perf_counter() - t_up, 3), "triangles": int(scene.n_tris), "built_on": "device" if ctx.built_on_device() else "host"}
    W, H = wl["width"], wl["height"]
    total_spp = args.spp * world
    s_begin, s_end = par.sample_range(total_spp, rank, world)  # shard by sample index
    desc = trb.render_desc(width=W, height=H, spp=total_spp, rr_depth=wl["rr_depth"], spp_begin=s_begin, spp_end=s_end,
                           wave_samples=args.wave)
    if args._visits:
        _, st = ctx.render_bdpt(desc, download=False)
        print(json.dumps({k: int(st[k]) for k in ("rays_closest", "rays_any", "node_visits", "tri_tests", "node_visits_any", "tri_tests_any", "samples",
                                                  "node_visits_bin", "tri_tests_bin", "node_visits_bin_any", "tri_tests_bin_any", "rays_deferred", "rays_deferred_any")}))
        return
    fb = torch.zeros(H, W, 3, device="cuda", dtype=torch.float32)
    flush = torch.empty(96 * 1024 * 1024, device="cuda", dtype=torch.float32)  # 384 MB > 126 MB L2
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        """device-resident step: render this rank's sample range into fb, then one reduce. returns (device ms, stats)"""
        flush.zero_()
        torch.cuda.synchronize()
        st = ctx.render_bdpt_device(desc, fb.data_ptr())  # kernels timed with CUDA events on the launching stream
        ms = st["ms_total"]
        if world > 1:
            ev0.record()
            par.reduce_framebuffer(fb, dst=0)  # the one collective of the render
            ev1.record()
            torch.cuda.synchronize()
            ms += ev0.elapsed_time(ev1)
        return ms, st

    for _ in range(args.warmup):
        step()
    clocks = ClockSampler(local)
    barrier()
    if rank == 0:
        clocks.start()
    tot_ms, agg = 0.0, {}
    for _ in range(args.steps):
        ms, st = step()
        tot_ms += ms
        for k, v in st.items():
            agg[k] = agg.get(k, 0) + v
    barrier()
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([tot_ms], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    tot_ms = float(t.item())
    samples_step = W * H * args.spp * world
    value = samples_step * args.steps / (tot_ms * 1e-3) / 1e6

    # ---- end to end through the host-buffer C ABI call: scene H2D (pinned) + render + frame buffer D2H ----
    e2e_steps = max(3, args.steps)
    ctx.refresh_scene()   # untimed: a device-built scene makes its pinned staging copy (the host side of the transfer) on the first refresh
    frame_t = torch.empty((H, W, 3), dtype=torch.float32, pin_memory=True)   # the caller's frame buffer, pinned: the D2H copy lands in it
    frame = frame_t.numpy()
    part = torch.zeros(H, W, 3, device="cuda", dtype=torch.float32) if world > 1 else None

    def e2e_step():
        ctx.refresh_scene()
        if world == 1:
            ctx.render_bdpt(desc, out=frame)
            return
        # N > 1: each rank renders its samples into a device frame, the partial frames meet on rank 0 over NCCL, rank 0 reads the sum back
        ctx.render_bdpt_device(desc, part.data_ptr())
        dist.reduce(part, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            frame_t.copy_(part)
        torch.cuda.synchronize()

    e2e_step()   # untimed: first use of the buffers and of the reduce of this shape
    gc.collect(); gc.disable()
    barrier()
    t0 = time.perf_counter()
    e2e_step_ms = []
    for _ in range(e2e_steps):
        ts = time.perf_counter()
        e2e_step()
        e2e_step_ms.append(round((time.perf_counter() - ts) * 1e3, 2))
    barrier()
    e2e_s = time.perf_counter() - t0
    gc.enable()
    te = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = samples_step * e2e_steps / float(te.item()) / 1e6

    # ---- configs[3] in the multi-GPU line: the 4K film (3840 x 2160, 99.5 MB frame buffer) partitioned by sample index over the ranks, one NCCL
    # reduce per step; a short pass (8 spp per GPU per step; the full 4096 spp are `--workload cbox4k --spp 512` on 8 GPUs) ----
    c4 = None
    if world > 1 and args.workload == "caustic" and not args.no_extras:
        w4 = WORKLOADS["cbox4k"]
        W4, H4, spp4 = w4["width"], w4["height"], 8
        b4, e4 = par.sample_range(spp4 * world, rank, world)
        d4 = trb.render_desc(width=W4, height=H4, spp=spp4 * world, rr_depth=w4["rr_depth"], spp_begin=b4, spp_end=e4, wave_samples=args.wave)
        fb4 = torch.zeros(H4, W4, 3, device="cuda", dtype=torch.float32)
        tot4 = 0.0
        for it in range(3):   # one warm-up, two timed
            flush.zero_(); barrier()
            st4 = ctx.render_bdpt_device(d4, fb4.data_ptr())
            ev0.record(); par.reduce_framebuffer(fb4, dst=0); ev1.record(); torch.cuda.synchronize()
            if it:
                tot4 += st4["ms_total"] + ev0.elapsed_time(ev1)
        t4 = torch.tensor([tot4], device="cuda", dtype=torch.float64)
        dist.all_reduce(t4, op=dist.ReduceOp.MAX)
        c4 = {"workload": w4["desc"], "value": W4 * H4 * spp4 * world * 2 / (float(t4.item()) * 1e-3) / 1e6, "unit": UNIT, "spp_per_gpu_per_step": spp4, "steps": 2, "warmup": 1,
              "ms_per_step": float(t4.item()) / 2, "reduce_bytes": W4 * H4 * 3 * 4, "image_mean": float(fb4.mean().item()) if rank == 0 else None,
              "timing": "device time of the render (CUDA events) + the NCCL reduce of the frame buffer, max over ranks"}
        del fb4
    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": tot_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
               "data": "synthetic",
               "config": {"workload": wl["desc"], "spp_per_gpu_per_step": args.spp, "samples_per_step": samples_step,
                          "partition": f"sample index, {world} rank(s), one NCCL reduce of the {W}x{H}x3 float frame buffer per step",
                          "l2": "384 MB buffer written between timed steps (L2 flush); wave buffers (>1 GB) exceed the 126 MB L2", "wave_samples": int(agg.get("wave_samples", 0) // max(args.steps, 1)),
                          "basis": "per sample; the reference arm (--impl reference) renders 1 spp per step, this arm --spp per step"},
               "clocks": clk, "gpu_launches": int(agg.get("launches", 0)),
               "scene_upload": upload_info,
               "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(ctx.scene_bytes()), "d2h_bytes_per_step": W * H * 3 * 4,
                       "steps": e2e_steps, "step_ms": e2e_step_ms, "timing": "host wall clock around trb_scene_refresh + trb_render_bdpt (frame buffer out into the caller's pinned host array; N > 1: trb_render_bdpt_device, NCCL reduce of the device frames, rank 0 reads the sum into its pinned array)",
                       "note": "the scene is re-sent from pinned memory EVERY step (the contract's letter: a step's inputs cross the bus inside the timed "
                               "region); a caller uploads a scene once and renders many frames - c3.e2e / c5.e2e are measured that way"},
               "mrays_per_s": {"closest": agg["rays_closest"] / (tot_ms * 1e-3) / 1e6, "any": agg["rays_any"] / (tot_ms * 1e-3) / 1e6,
                               "rays_per_sample": (agg["rays_closest"] + agg["rays_any"]) / max(agg["samples"], 1)},
               "stage_ms_per_step": {k: agg[k] / args.steps for k in ("ms_trace_closest", "ms_trace_any", "ms_logic")},
               "stage_ms_note": "kernel time summed per kind; shadow traversal runs on a second stream, so the kinds overlap and add up to more than ms_per_step"}
        # ---- roofline of the dominant kernel: per-kernel time from an extra pass with the two streams serialised
        # (TRB_FLAG_NO_OVERLAP), so that the kernel's duration is its own and not a share of an overlapped interval ----
        ser = serial_pass(trb, torch, ctx, fb, flush, wl, W, H, total_spp, s_begin, s_end, args.wave)
        out["stage_ms_serial"] = {k: ser[k] for k in ("ms_total", "ms_trace_closest", "ms_trace_any", "ms_logic")}
        vis = visits_subprocess(args.workload, args.spp, args.wave) if (world == 1 and not args.no_visits) else None
        out["roofline"] = roofline_block(args.workload, wl, ser, vis, W * H * args.spp, args.wave, ctx.scene_bytes())
        # ---- the other single-GPU configurations the targets are judged on, each a short pass of its own (N = 1, default workload only) ----
        if world == 1 and args.workload == "caustic" and not args.no_extras:
            for key, name, spp in (("c3", "hardlight", 8), ("c5", "tess", 16)):   # (16 spp = one 2^24-sample wave at 1024^2)
                try:
                    out[key] = extra_block(trb, torch, name, spp, fb, flush, args)
                except Exception as e:  # noqa: BLE001
                    out[key] = {"error": str(e)}
            if "roofline" in out.get("c5", {}):
                out["roofline_c5"] = out["c5"]["roofline"]
        # ---- the reference's CPU renderer on this box's host cores (a reported baseline, not the target) ----
        if world == 1 and not args.no_cpu_baseline:
            try:
                with quiet_stdout():
                    run, info = cpu_reference_step_factory(wl)
                    n, s, cpu_img = run(want_image=True)
                fw, fh = info.pop("film")
                out["cpu_baseline"] = dict(info, value=n / s / 1e6, unit=UNIT)
                if wl["scene"] is not None:  # (the tess workload times the CPU on a smaller instance of the scene: no common image)
                    out["equal_time"] = equal_time_relmse(trb, ctx, wl, fw, fh, cpu_img, s, value * 1e6)
                    out["cpu_ceiling"] = cpu_port_all_cores(trb, wl, fw, fh)
            except Exception as e:  # noqa: BLE001
                out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "unavailable", "sample": str(e)}
        if c4 is not None:
            out["c4"] = c4
        if world > 1 and not args.no_extras:
            # the same frame through ONE process (trb_render_bdpt_multi: what the C++ API / CLI use for gpus = N), next to the torchrun value
            out["inproc_multi"] = inproc_subprocess(args, world)
        print(json.dumps(out))
    if world > 1:
        if cpu_group is not None:
            dist.barrier(group=cpu_group)   # the other ranks wait on the CPU: an NCCL barrier is a spinning kernel on their GPUs, which the
                                            # in-process run of rank 0's child would time-slice against
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a._inproc > 0:
        with quiet_stdout():
            import io
            buf = io.StringIO(); real = sys.stdout; sys.stdout = buf
            try:
                run_inproc_child(a)
            finally:
                sys.stdout = real
        sys.stdout.write(buf.getvalue()); sys.stdout.flush()
    elif a.impl == "reference":
        with quiet_stdout():
            import io
            buf = io.StringIO()
            real = sys.stdout
            sys.stdout = buf
            try:
                run_reference_arm(a)
            finally:
                sys.stdout = real
        sys.stdout.write(buf.getvalue())
        sys.stdout.flush()
    else:
        run_ours(a)
