#!/usr/bin/env python
"""Small run of every kernel family, meant to be executed under compute-sanitizer on the GPU box:

    compute-sanitizer --tool memcheck  --error-exitcode 9 python tools/sanitize_run.py
    compute-sanitizer --tool racecheck --error-exitcode 9 python tools/sanitize_run.py
    compute-sanitizer --tool initcheck --error-exitcode 9 python tools/sanitize_run.py

Sizes are tiny (64x64, 2 spp, a 20 k-triangle atrium) because the tools slow the kernels down 10-100x.  No oracle, no
checks of the values: the tools' reports are the result.  (On this GPU pool compute-sanitizer is closed — rounds 1 and 2 — so
the plain run is what there is: it walks every kernel family and entry point once and was clean.)"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cudapathtracer_b200 import api, scenes

W = int(os.environ.get("W", "64"))
SPP = int(os.environ.get("SPP", "2"))


def frame(ctx, what, **opts):
    ctx.set_options(**opts)
    ctx.reset_accum()
    ctx.render(0, SPP)
    img = ctx.read_accum(True)
    print(f"{what}: mean {float(img.mean()):.4f}", flush=True)
    return img


def run(scene, name, spheres=False):
    s = api.SceneHandle()
    s.add_scene(scene)
    v, t, m, l = s.arrays()
    n2, d2 = s.build_bvh2()
    rank = api.bvh2_leaf_order(n2, len(t))
    bvh = s.build_bvh(tie_rank=rank)
    cam = api.make_camera(W, W, **scene.camera)
    ctx = api.Context(0)
    ctx.set_scene(s)
    ctx.set_bvh(bvh)
    ctx.set_camera(cam)
    ctx.set_options(max_depth=5, seed=1234, samples_per_pass=0)
    ctx.primary_hits()
    frame(ctx, f"{name} pt wide", integrator=api.INTEGRATOR_PT, traversal=api.TRAVERSAL_WIDE)
    frame(ctx, f"{name} pt wide stats", collect_stats=1)
    ctx.read_heatmap()
    ctx.set_options(collect_stats=0)
    frame(ctx, f"{name} pt shared primary", share_primary_hits=1)
    ctx.set_options(share_primary_hits=0)
    if not spheres:
        ctx.set_bvh2(n2, d2)
        frame(ctx, f"{name} pt strict", traversal=api.TRAVERSAL_WIDE_STRICT, record_retraces=1)
        ctx.read_retrace_counts()
        ctx.set_options(record_retraces=0)
        frame(ctx, f"{name} pt bvh2", traversal=api.TRAVERSAL_REF_BVH2)
        frame(ctx, f"{name} pt watertight", traversal=api.TRAVERSAL_WIDE, tri_test=api.TRI_WATERTIGHT)
        ctx.set_options(tri_test=api.TRI_REFEXACT)
    if not spheres:   # (the BDPT integrator has no sphere support)
        frame(ctx, f"{name} bdpt wide", integrator=api.INTEGRATOR_BDPT, traversal=api.TRAVERSAL_WIDE)
        ctx.set_options(integrator=api.INTEGRATOR_PT)
    nt = ctx.tile_count()
    ctx.set_tile_window(nt // 2, nt - nt // 2)
    frame(ctx, f"{name} pt tile window")
    ctx.set_tile_window(0, 0)
    ctx.read_image_rgb8(mirror_x=True)
    rays = ctx.camera_rays()
    ctx.trace_rays(rays[:1000])
    ctx.trace_rays_noalloc(rays[:1000]); ctx.trace_rays_noalloc(rays[:77])
    g = api.Group(devices=[0])   # the group path with one member: broadcast header, workers, in-place reduce skipped
    g.set_scene(s); g.set_bvh(bvh); g.set_camera(cam); g.set_options(max_depth=5, seed=1234)
    g.reset_accum(); g.render(0, SPP); g.read_accum(True); g.read_image_rgb8(); g.close()
    info = ctx.build_bvh_device(tie_rank=rank)
    print(f"{name} device bvh: {info}", flush=True)
    frame(ctx, f"{name} pt wide, device bvh", traversal=api.TRAVERSAL_WIDE)
    ctx.read_bvh()
    ctx.close()


run(scenes.cornell_box(), "cornell")
run(scenes.cornell_box(extra_lights=8, spheres=True, specular=True), "cornell+spheres+mirrors", spheres=True)
run(scenes.atrium(20000), "atrium20k")
print("sanitize_run done")
