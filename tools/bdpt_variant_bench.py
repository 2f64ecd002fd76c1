"""Developer tool (GPU box): one line per build of libcupt.so — config 2 through the reference's live integrator (BDPT): frame time by stage, image checksum.

    CUPT_LIB=variants/libcupt_x.so python tools/bdpt_variant_bench.py [label]"""
import os, sys, time
sys.path.insert(0, os.getcwd())
from cudapathtracer_b200 import api, scenes
sc = scenes.atrium(262144)
s = api.SceneHandle(); s.add_scene(sc)
b = s.build_bvh()
ctx = api.Context(0)
ctx.set_scene(s); ctx.set_bvh(b); ctx.set_camera(api.make_camera(1024, 1024, **sc.camera))
ctx.set_options(max_depth=5, seed=1234, traversal=api.TRAVERSAL_WIDE, integrator=api.INTEGRATOR_BDPT, time_stages=1, collect_stats=0)
ctx.reset_accum(); ctx.render(0, 64); ctx.sync()
best, st = 1e9, None
for _ in range(3):
    ctx.reset_accum(); ctx.reset_stats(); t0 = time.perf_counter(); ctx.render(0, 64); ctx.sync(); dt = time.perf_counter() - t0
    if dt < best: best, st = dt, ctx.stats()
import hashlib
print(f"{sys.argv[1]:8s} bdpt frame {best*1e3:7.2f} ms gen {st['ms_raygen']:.2f} extend {st['ms_extend']:.2f} mid+connect {st['ms_shade']:.2f} resolve+accum {st['ms_accum']:.2f} image {hashlib.sha1(ctx.read_accum(False).tobytes()).hexdigest()[:10]}", flush=True)
