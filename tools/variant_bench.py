"""Developer tool (GPU box): one line per build of libcupt.so — config 2 frame time by stage, rays, image checksum.

    CUPT_LIB=variants/libcupt_x.so python tools/variant_bench.py [label]

Used with tools/build_variants.sh for same-box A/B runs of compile-time choices (the product reads no environment)."""
import hashlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudapathtracer_b200 import api, scenes

def main():
    label = sys.argv[1] if len(sys.argv) > 1 else os.path.basename(api.LIB_PATH)
    spp = int(os.environ.get("SPP", "64"))
    sc = scenes.atrium(262144)
    s = api.SceneHandle(); s.add_scene(sc)
    n2, d2 = s.build_bvh2()
    b = s.build_bvh(tie_rank=api.bvh2_leaf_order(n2, s.counts()[1]), reinsert_passes=int(os.environ.get("REINSERT", "0")), sah_bins=int(os.environ.get("BINS", "0")))
    ctx = api.Context(0)
    ctx.set_scene(s); ctx.set_bvh(b); ctx.set_camera(api.make_camera(1024, 1024, **sc.camera))
    ctx.set_options(max_depth=5, seed=1234, traversal=api.TRAVERSAL_WIDE, time_stages=1, collect_stats=0)
    ctx.reset_accum(); ctx.render(0, spp); ctx.sync()
    best, st = 1e9, None
    for _ in range(int(os.environ.get("REPS", "4"))):
        ctx.reset_accum(); ctx.reset_stats()
        t0 = time.perf_counter(); ctx.render(0, spp); ctx.sync(); dt = time.perf_counter() - t0
        if dt < best: best, st = dt, ctx.stats()
    h = hashlib.sha1(ctx.read_accum(False).tobytes()).hexdigest()[:12]
    print(f"{label:28s} frame {best*1e3:7.2f} ms  extend {st['ms_extend']:6.2f}  shade {st['ms_shade']:5.2f}  raygen {st['ms_raygen']:4.2f}  accum {st['ms_accum']:5.2f}  "
          f"{st['rays']/best/1e6:7.1f} Mrays/s  image {h}", flush=True)
    ctx.close()

if __name__ == "__main__":
    main()
