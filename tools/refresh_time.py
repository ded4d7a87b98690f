import importlib, sys, time
sys.path.insert(0, '.'); sys.path.insert(0, 'tools')
import gen_tess_scene
trb = importlib.import_module("bidirectional-path-tracing_b200")
sc = gen_tess_scene.make_scene(10_000_000, seed=1)
for mode in ("host", "device"):
    ctx = trb.Context(0); ctx.upload(sc, build=mode)
    ts = []
    for i in range(5):
        t0 = time.perf_counter(); ctx.refresh_scene(); ts.append(time.perf_counter() - t0)
    print(mode, [round(t * 1e3, 1) for t in ts], flush=True)
    ctx.close()
