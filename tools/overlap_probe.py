"""Developer tool (GPU box): would the extend stage of one half of the frame overlap the shade stage of the other half?
Two contexts on ONE GPU render samples [0, 32) and [32, 64) of config 2's frame concurrently (two streams), against one context
rendering all 64.  CUPT_LIB selects a build (tools/build_variants.sh with -DCUPT_EXTEND_MAX_BLOCKS=k leaves room on the SMs)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudapathtracer_b200 import api, scenes

label = sys.argv[1] if len(sys.argv) > 1 else os.path.basename(api.LIB_PATH)
sc = scenes.atrium(262144)
s = api.SceneHandle(); s.add_scene(sc)
n2, d2 = s.build_bvh2()
b = s.build_bvh(tie_rank=api.bvh2_leaf_order(n2, s.counts()[1]))
cam = api.make_camera(1024, 1024, **sc.camera)
ctxs = [api.Context(0) for _ in range(2)]
for c in ctxs:
    c.set_scene(s); c.set_bvh(b); c.set_camera(cam)
    c.set_options(max_depth=5, seed=1234, traversal=api.TRAVERSAL_WIDE)

def run(parts):
    best = 1e9
    for _ in range(5):
        for c in ctxs: c.reset_accum()
        for c in ctxs: c.sync()
        t0 = time.perf_counter()
        for c, (first, n) in zip(ctxs, parts):
            if n: c.render(first, n)
        for c in ctxs: c.sync()
        best = min(best, time.perf_counter() - t0)
    return best * 1e3

one = run([(0, 64), (0, 0)])
two = run([(0, 32), (32, 32)])
four = None
print(f"{label:24s} one context 64 spp {one:7.2f} ms   two contexts 32 + 32 spp {two:7.2f} ms   ratio {two / one:.3f}", flush=True)
for c in ctxs: c.close()
