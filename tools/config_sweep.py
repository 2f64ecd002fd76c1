#!/usr/bin/env python
"""BASELINE.json configs 1, 3 (one GPU's share) and 5 as measurement lines; config 2 is bench.py, config 4 is
tools/soup_sweep.py.  GPU box tool: one JSON line per config into gpurun_out/configs.jsonl.

  config 1  Cornell + spheres 256x256, 16 spp, depth 5 (also timed: the reference integrator has no sphere code, so
            the CPU figure is the oracle restatement)
  config 3  atrium 3840x2160, SPP3 samples (default 128 = one GPU's share of 1024 spp on 8 GPUs)
  config 5  Cornell + 8 extra triangle lights + mirror tall box / back wall, 1024x1024, 64 spp, depth 16: wavefront path
            tracer, alive paths per depth; and the rebuilt kernel.cu (its BDPT megakernel, no mirrors) next to this
            library's BDPT on the same diffuse scene
"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cudapathtracer_b200 import api, scenes

OUT = open(os.path.join(ROOT, "gpurun_out", "configs.jsonl"), "a")
ctx = api.Context(0)


def timed(ctx, spp, reps=3):
    ctx.set_options(collect_stats=0, time_stages=1)
    ctx.reset_accum(); ctx.render(0, spp); ctx.sync()
    best, st = 1e9, None
    for _ in range(reps):
        ctx.reset_accum(); ctx.reset_stats()
        t0 = time.perf_counter(); ctx.render(0, spp); ctx.sync(); dt = time.perf_counter() - t0
        if dt < best: best, st = dt, ctx.stats()
    return best, st


def emit(rec):
    print(json.dumps(rec), flush=True); OUT.write(json.dumps(rec) + "\n"); OUT.flush()


def setup(sc, w, h, **opts):
    s = api.SceneHandle(); s.add_scene(sc)
    cam = api.make_camera(w, h, **sc.camera)
    ctx.set_scene(s); ctx.set_camera(cam)
    info = ctx.build_bvh_device()
    ctx.set_options(traversal=api.TRAVERSAL_WIDE, seed=1234, integrator=api.INTEGRATOR_PT, **opts)
    return s, cam, info


which = os.environ.get("CONFIGS", "1,3,5").split(",")
if "1" in which:
    s, cam, info = setup(scenes.cornell_box(spheres=True), 256, 256, max_depth=5)
    dt, st = timed(ctx, 16)
    emit({"config": 1, "scene": "cornell + 2 spheres (sphere.h)", "image": "256x256", "spp": 16, "max_depth": 5, "ms": round(dt * 1e3, 3),
          "mrays_per_s": round(st["rays"] / dt / 1e6, 1), "msamples_per_s": round(256 * 256 * 16 / dt / 1e6, 1), "rays_per_sample": round(st["rays"] / (256 * 256 * 16), 3),
          "note": "1 Mi paths: launch-latency bound (12 launches per frame)"})
if "3" in which:
    spp3 = int(os.environ.get("SPP3", "128"))
    s, cam, info = setup(scenes.atrium(262144, seed=1), 3840, 2160, max_depth=5)
    dt, st = timed(ctx, spp3, reps=2)
    emit({"config": 3, "scene": "atrium 262144 tris", "image": "3840x2160", "spp_this_gpu": spp3, "max_depth": 5, "ms": round(dt * 1e3, 1),
          "mrays_per_s": round(st["rays"] / dt / 1e6, 1), "msamples_per_s": round(3840 * 2160 * spp3 / dt / 1e6, 1),
          "stage_ms": {k: round(st["ms_" + k], 1) for k in ("raygen", "extend", "shade", "accum")},
          "note": "one GPU's sample range of the 1024-spp frame; the N-GPU job adds one NCCL reduce of 99.5 MB of f32 sums"})
if "5" in which:
    W = 1024
    s, cam, info = setup(scenes.cornell_box(extra_lights=8, specular=True), W, W, max_depth=16)
    dt, st = timed(ctx, 64)
    alive = []
    prev = 0
    for d in range(1, 17):   # rays traced with max_depth = d minus those with d - 1 = paths alive at depth d
        ctx.set_options(max_depth=d, collect_stats=1); ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 4)
        r = ctx.stats()["rays"]; alive.append(round((r - prev) / (W * W * 4), 4)); prev = r
    ctx.set_options(max_depth=16, collect_stats=0)
    rec = {"config": 5, "scene": "cornell + 8 triangle lights, mirror tall box and back wall", "image": "1024x1024", "spp": 64, "max_depth": 16,
           "ms": round(dt * 1e3, 2), "mrays_per_s": round(st["rays"] / dt / 1e6, 1), "msamples_per_s": round(W * W * 64 / dt / 1e6, 1),
           "rays_per_sample": round(st["rays"] / (W * W * 64), 3), "alive_fraction_by_depth": alive,
           "stage_ms": {k: round(st["ms_" + k], 2) for k in ("raygen", "extend", "shade", "accum")}}
    # megakernel vs wavefront on the same DIFFUSE scene with the reference's live integrator
    s2, cam2, _ = setup(scenes.cornell_box(extra_lights=8), W, W, max_depth=5)
    ctx.set_options(integrator=api.INTEGRATOR_BDPT)
    dt2, st2 = timed(ctx, 64)
    rec["bdpt_wavefront"] = {"ms": round(dt2 * 1e3, 2), "mrays_per_s": round(st2["rays"] / dt2 / 1e6, 1), "msamples_per_s": round(W * W * 64 / dt2 / 1e6, 1)}
    try:
        import legacybind
        if legacybind.available() and legacybind.image_dim() == W:
            v, t, m, l = s2.arrays()
            r = legacybind.render(v, t, m, l, s2.total_light_area, cam2, 64)
            rec["bdpt_megakernel_legacy"] = {"ms": round(r["ms"], 2), "mrays_per_s": round(r["rays"] / (r["ms"] * 1e-3) / 1e6, 1),
                                             "msamples_per_s": round(W * W * 64 / (r["ms"] * 1e-3) / 1e6, 1)}
    except Exception as e:
        rec["bdpt_megakernel_legacy"] = {"unavailable": str(e)[:160]}
    emit(rec)
