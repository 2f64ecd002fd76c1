#!/usr/bin/env python
"""BDPT frame time on config 2 against the pass size (samples of every pixel in flight at once)."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from cudapathtracer_b200 import api
sc, s, cam = bench.build_scene()
ctx = api.Context(0)
ctx.set_scene(s); ctx.set_camera(cam); ctx.set_bvh(s.build_bvh())
for spass in [int(x) for x in os.environ.get("SPASS", "8,16,32,64").split(",")]:
    ctx.set_options(max_depth=5, seed=1234, traversal=api.TRAVERSAL_WIDE, integrator=api.INTEGRATOR_BDPT, samples_per_pass=spass, time_stages=1)
    ctx.reset_accum(); ctx.render(0, 64); ctx.sync()
    best = 1e9
    for _ in range(3):
        ctx.reset_accum(); ctx.reset_stats()
        t0 = time.perf_counter(); ctx.render(0, 64); ctx.sync(); dt = time.perf_counter() - t0
        if dt < best: best, st = dt, ctx.stats()
    print(json.dumps({"samples_per_pass": spass, "state_gb": round(spass * 1024 * 1024 * 288 / 1e9, 1), "ms": round(best * 1e3, 2), "ms_extend": round(st["ms_extend"], 2), "mrays": round(st["rays"] / best / 1e6)}), flush=True)
