"""Developer tool (GPU box): time the headline step for several extend tunings.  Not a test."""
import os, sys, time, itertools
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cudapathtracer_b200 import api, scenes

def main():
    W = 1024; SPP = int(os.environ.get("SPP", "32"))
    sc = scenes.atrium(262144)
    s = api.SceneHandle(); s.add_scene(sc)
    b = s.build_bvh(sah_bins=int(os.environ.get("BINS", "0")), max_leaf_tris=int(os.environ.get("LEAF", "0")))
    print("bvh", b.info(), flush=True)
    cam = api.make_camera(W, W, **sc.camera)
    ctx = api.Context(0)
    ctx.set_scene(s); ctx.set_bvh(b); ctx.set_camera(cam)
    combos = [tuple(int(x) for x in c.split(":")) for c in os.environ.get("COMBOS", "24:8,24:1,16:8,28:8,24:4,24:12,24:16,20:6,31:8,12:1").split(",")]
    ref_img = None
    for ft, tt in combos:
        ctx.set_options(max_depth=5, seed=1234, traversal=api.TRAVERSAL_WIDE, time_stages=1, collect_stats=0, fetch_threshold=ft, defer_pairs=tt, samples_per_pass=int(os.environ.get("SPASS", "0")))
        ctx.reset_accum(); ctx.render(0, 4); ctx.sync()
        best = 1e9
        for rep in range(3):
            ctx.reset_accum(); ctx.reset_stats()
            t0 = time.perf_counter(); ctx.render(0, SPP); ctx.sync(); dt = time.perf_counter() - t0
            best = min(best, dt)
            st = ctx.stats()
        img = ctx.read_accum(False)
        same = "ref" if ref_img is None else ("same" if img.tobytes() == ref_img.tobytes() else f"DIFF max {np.abs(img-ref_img).max():.3g}")
        if ref_img is None: ref_img = img
        ctx.set_options(collect_stats=1, time_stages=0)
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 2); st2 = ctx.stats()
        print(f"fetch {ft:2d} tri {tt:2d}: {best*1e3:7.2f} ms  {st['rays']/best/1e6:8.1f} Mrays/s  extend {st['ms_extend']:.1f} shade {st['ms_shade']:.1f} "
              f"nodes/ray {st2['nodes_visited']/st2['rays']:.2f} tris/ray {st2['tris_tested']/st2['rays']:.2f}  image {same}", flush=True)
    ctx.close()

if __name__ == "__main__":
    main()
