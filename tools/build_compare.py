#!/usr/bin/env python
"""Host (binned SAH, OpenMP) vs device (Morton + PLOC) wide-BVH builders on config 2: build time, tree statistics and
the render time of the same frame through each tree.  GPU box tool; prints one JSON line per builder."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from cudapathtracer_b200 import api

sc, s, cam = bench.build_scene()
ctx = api.Context(0)
ctx.set_scene(s); ctx.set_camera(cam)
for which in ["host"] + ["device" + r for r in os.environ.get("RADII", "1,2,4,8").split(",")]:
    if which == "host":
        t0 = time.perf_counter(); b = s.build_bvh(); ctx.set_bvh(b); wall = time.perf_counter() - t0
        info = b.info()
    else:
        R = int(which[6:])
        ctx.build_bvh_device(radius=R)
        t0 = time.perf_counter(); info = ctx.build_bvh_device(radius=R); wall = time.perf_counter() - t0
    rec = {"builder": which, "build_wall_ms": round(wall * 1e3, 2), **{k: (round(v, 3) if isinstance(v, float) else v) for k, v in info.items()}}
    for integ, name in ((api.INTEGRATOR_PT, "pt"), (api.INTEGRATOR_BDPT, "bdpt")):
        ctx.set_options(max_depth=bench.DEPTH, seed=bench.SEED, traversal=api.TRAVERSAL_WIDE, integrator=integ, collect_stats=1, time_stages=0)
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 4); st = ctx.stats()
        rec[name + "_nodes_per_ray"] = round(st["nodes_visited"] / st["rays"], 2); rec[name + "_tris_per_ray"] = round(st["tris_tested"] / st["rays"], 2)
        ctx.set_options(collect_stats=0, time_stages=1)
        ctx.reset_accum(); ctx.render(0, bench.SPP); ctx.sync()
        best = 1e9
        for _ in range(3):
            ctx.reset_accum(); ctx.reset_stats()
            t0 = time.perf_counter(); ctx.render(0, bench.SPP); ctx.sync(); dt = time.perf_counter() - t0
            if dt < best: best, stb = dt, ctx.stats()
        rec[name + "_ms"] = round(best * 1e3, 2); rec[name + "_ms_extend"] = round(stb["ms_extend"], 2); rec[name + "_mrays"] = round(stb["rays"] / best / 1e6, 1)
    print(json.dumps(rec), flush=True)
