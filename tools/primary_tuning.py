"""Developer tool (GPU box): the depth-0 (primary ray) extend launch of config 2 alone (max_depth 1) for several extend tunings."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudapathtracer_b200 import api, scenes

sc = scenes.atrium(262144)
s = api.SceneHandle(); s.add_scene(sc)
b = s.build_bvh()
ctx = api.Context(0)
ctx.set_scene(s); ctx.set_bvh(b); ctx.set_camera(api.make_camera(1024, 1024, **sc.camera))
for ft, dp in [(24, 20), (24, 1), (24, 4), (24, 8), (24, 12), (24, 32), (28, 20), (31, 20), (16, 20), (31, 1)]:
    ctx.set_options(max_depth=1, seed=1234, traversal=api.TRAVERSAL_WIDE, time_stages=1, collect_stats=0, fetch_threshold=ft, defer_pairs=dp)
    ctx.reset_accum(); ctx.render(0, 64); ctx.sync()
    best = 1e9
    for _ in range(3):
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 64); ctx.sync()
        best = min(best, ctx.stats()["ms_extend"])
    ctx.set_options(collect_stats=1, time_stages=0)
    ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 4); st = ctx.stats()
    print(f"fetch {ft:2d} defer {dp:2d}: primary extend {best:6.2f} ms  nodes/ray {st['nodes_visited']/st['rays']:.2f} tris/ray {st['tris_tested']/st['rays']:.2f}", flush=True)
