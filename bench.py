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
    ap.add_argument("--no-visits", action="store_true", help="skip the counters-build pass (A/B runs: no roofline figure)")
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
def cpu_reference_step_factory(workload: dict, budget_s: float = 8.0):
    """Returns (run_step, info): run_step() renders one bounded sample of the workload with the reference's own
    threaded renderer and returns (samples, seconds).  Film is the workload's unless a calibration run says one
    step would take much longer than budget_s, then it is halved (same scene, camera, rrDepth)."""
    import refh
    import scene_to_obj
    trb = importlib.import_module(PKG)
    scene = load_scene(trb, workload, cpu=True)
    tmp = Path(tempfile.mkdtemp(prefix="trb_ref_"))
    obj = scene_to_obj.export_verified(scene, tmp / "scene.obj")
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
    run, info = cpu_reference_step_factory(wl)
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
                      "config": {"workload": wl["desc"], "spp_per_step": 1, "timing": "reference's own wall clock around Renderer::render"},
                      "cpu_baseline": info, "gpu_launches": 0,
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


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


def visits_subprocess(args) -> dict | None:
    """node / triangle visit counts per ray class from the counters build (untimed, one step)."""
    cnt = ROOT / PKG / "libtrb_counters.so"
    if not cnt.exists():
        return None
    env = dict(os.environ, TRB_LIB=str(cnt))
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    try:
        r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--_visits", "--workload", args.workload, "--spp", str(min(args.spp, 4)),
                            "--wave", str(args.wave)], env=env, capture_output=True, text=True, timeout=600)
        return json.loads(r.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        sys.stderr.write(f"visit counting failed: {e}\n")
        return None


def run_ours(args):
    import torch
    import torch.distributed as dist
    trb = importlib.import_module(PKG)
    par = importlib.import_module(PKG + ".parallel")
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    wl = WORKLOADS[args.workload]
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
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
        print(json.dumps({k: int(st[k]) for k in ("rays_closest", "rays_any", "node_visits", "tri_tests", "node_visits_any", "tri_tests_any", "samples")}))
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
    e2e_steps = max(2, min(args.steps, 3))
    ctx.refresh_scene()   # untimed: a device-built scene makes its pinned staging copy (the host side of the transfer) on the first refresh
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        ctx.refresh_scene()
        img, _ = ctx.render_bdpt(desc, download=True)
        if world > 1:  # the partial frame buffers meet on rank 0 (host result is rank 0's sum)
            part = torch.from_numpy(img).cuda()
            dist.reduce(part, dst=0, op=dist.ReduceOp.SUM)
            img = part.cpu().numpy()
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = samples_step * e2e_steps / float(te.item()) / 1e6

    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": tot_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
               "data": "synthetic",
               "config": {"workload": wl["desc"], "spp_per_gpu_per_step": args.spp, "samples_per_step": samples_step,
                          "partition": f"sample index, {world} rank(s), one NCCL reduce of the {W}x{H}x3 float frame buffer per step",
                          "l2": "384 MB buffer written between timed steps (L2 flush); wave buffers (>1 GB) exceed the 126 MB L2"},
               "clocks": clk, "gpu_launches": int(agg.get("launches", 0)),
               "scene_upload": upload_info,
               "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(ctx.scene_bytes()), "d2h_bytes_per_step": W * H * 3 * 4,
                       "steps": e2e_steps, "timing": "host wall clock around trb_scene_refresh + trb_render_bdpt (host frame buffer out)"},
               "mrays_per_s": {"closest": agg["rays_closest"] / (tot_ms * 1e-3) / 1e6, "any": agg["rays_any"] / (tot_ms * 1e-3) / 1e6,
                               "rays_per_sample": (agg["rays_closest"] + agg["rays_any"]) / max(agg["samples"], 1)},
               "stage_ms_per_step": {k: agg[k] / args.steps for k in ("ms_trace_closest", "ms_trace_any", "ms_logic")},
               "stage_ms_note": "kernel time summed per kind; shadow traversal runs on a second stream, so the kinds overlap and add up to more than ms_per_step"}
        # ---- roofline of the dominant kernel: per-kernel time from an extra pass with the two streams serialised
        # (TRB_FLAG_NO_OVERLAP), so that the kernel's duration is its own and not a share of an overlapped interval ----
        desc_serial = trb.render_desc(width=W, height=H, spp=total_spp, rr_depth=wl["rr_depth"], spp_begin=s_begin, spp_end=s_end,
                                      wave_samples=args.wave, flags=trb.TRB_FLAG_NO_OVERLAP)
        ser = {}
        for _ in range(2):
            flush.zero_(); torch.cuda.synchronize()
            ser = ctx.render_bdpt_device(desc_serial, fb.data_ptr())
        out["stage_ms_serial"] = {k: ser[k] for k in ("ms_total", "ms_trace_closest", "ms_trace_any", "ms_logic")}
        agg_s = {k: ser[k] * args.steps for k in ser}  # same units as agg (sums over args.steps steps)
        peak, how = measured_peak()
        vis = visits_subprocess(args) if (world == 1 and not args.no_visits) else None
        any_dom = agg_s["ms_trace_any"] >= agg_s["ms_trace_closest"]
        dom = "k_trace_coop<1>" if any_dom else "k_trace_coop<0>"   # <1>: occlusion + resolve, <0>: closest hit
        K = wl["rr_depth"] - 1
        waves = -(-(W * H * args.spp) // (args.wave if args.wave > 0 else 1 << 22))
        dom_launches = (2 * K if any_dom else 2 * K) * waves * args.steps  # per wave: K light + K eye (any) / 1 + K + (K-1) (closest)
        scene_note = ("acceleration structure ~0.8 GB: HBM/L2-resident, the HBM-bound case" if args.workload == "tess" else
                      "this scene's BVH (~2.5 MB nodes+triangles) is L1/L2-resident, so the figure measures issue/L1/L2-side efficiency "
                      "against the HBM peak (ncu: DRAM traffic is a few % of the algorithmic bytes)")
        roof = {"bound": "hbm", "kernel": dom, "achieved": None, "peak": peak, "unit": "GB/s", "frac": None, "traffic": None, "peak_source": how}
        if vis:
            tag = "_any" if any_dom else ""
            nrays_v = max(vis["rays_any" if any_dom else "rays_closest"], 1)
            per_ray = 32 + 16 + 64 * vis["node_visits" + tag] / nrays_v + 48 * vis["tri_tests" + tag] / nrays_v
            nrays, ms = (agg_s["rays_any"], agg_s["ms_trace_any"]) if any_dom else (agg_s["rays_closest"], agg_s["ms_trace_closest"])
            achieved = per_ray * nrays / (ms * 1e-3) / 1e9
            roof.update(achieved=achieved, frac=achieved / peak, bytes_per_ray=per_ray, rays_per_launch=nrays / dom_launches,
                        bytes_per_launch=per_ray * nrays / dom_launches, ms_per_launch=ms / dom_launches,
                        nodes_per_ray=vis["node_visits" + tag] / nrays_v, tris_per_ray=vis["tri_tests" + tag] / nrays_v)
        tf = ROOT / "profiles" / "dram_traffic.json"
        if tf.exists():
            try:
                t = json.loads(tf.read_text())[args.workload]
                roof["traffic"] = t["kernels"][dom]["dram_bytes_per_launch"]
                roof["traffic_source"] = f"profiles/dram_traffic.json ({t['source']}: ncu dram__bytes_read+write per launch, {t['spp']} spp)"
            except Exception:
                pass
        roof["note"] = ("algorithmic bytes = rays x (32 B ray + 16 B result/pending + 64 B x nodes visited + 48 B x triangles tested), "
                        "visit counts from the -DTRB_COUNTERS twin library on the same workload; " + scene_note)
        out["roofline"] = roof
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
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
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
