#!/usr/bin/env python
"""Wall-clock breakdown of bench.py's e2e step (scene + BVH upload, render, image read-back) per integrator."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from cudapathtracer_b200 import api

sc, s, cam = bench.build_scene()
bvh = s.build_bvh()
ctx = api.Context(0)
img = np.empty((bench.W * bench.H, 3), dtype=np.float32)
for integ in (api.INTEGRATOR_PT, api.INTEGRATOR_BDPT):
    ctx.set_scene(s); ctx.set_bvh(bvh); ctx.set_camera(cam)
    ctx.set_options(max_depth=bench.DEPTH, seed=bench.SEED, integrator=integ, time_stages=int(os.environ.get("STAGES", "1")))
    for it in range(3):
        t = [time.perf_counter()]
        ctx.set_scene(s); t.append(time.perf_counter())
        ctx.set_bvh(bvh); ctx.set_camera(cam); t.append(time.perf_counter())
        ctx.reset_accum(); ctx.render(0, bench.SPP); t.append(time.perf_counter())
        ctx.sync(); t.append(time.perf_counter())
        ctx.read_accum(False, out=img); t.append(time.perf_counter())
        print(integ, it, " ".join(f"{n}={1e3 * (b - a):.1f}" for n, a, b in zip(("set_scene", "set_bvh", "render_call", "sync", "read"), t, t[1:])), flush=True)
