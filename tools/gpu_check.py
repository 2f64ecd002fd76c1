"""Developer diagnostic (GPU box): parity of the CUDA path against the oracle + first timings.  Not a test."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cudapathtracer_b200 import api, scenes
import oraclebind

def setup(sc, W, H, with_rank=True):
    s = api.SceneHandle(); s.add_scene(sc)
    v, t, m, l = s.arrays()
    n2, d2 = s.build_bvh2()
    rank = api.bvh2_leaf_order(n2, len(t)) if with_rank else None
    b = s.build_bvh(tie_rank=rank)
    osc = oraclebind.OracleScene(v, t, m, l, s.total_light_area, n2)
    cam = api.make_camera(W, H, **sc.camera)
    ctx = api.Context(0)
    ctx.set_scene(s); ctx.set_bvh(b); ctx.set_bvh2(n2, d2); ctx.set_camera(cam)
    return s, b, osc, cam, ctx, rank

def cmp_hits(tag, ids, ts, oi, ot):
    bad = np.nonzero((ids != oi) | (ts.view(np.uint32) != ot.view(np.uint32)))[0]
    print(f"  {tag}: {len(ids)} rays, mismatches {len(bad)}", flush=True)
    return bad

def main():
    for name, sc, W in (("cornell", scenes.cornell_box(), 256), ("atrium", scenes.atrium(), 1024)):
        print(f"== {name} {W}x{W}", flush=True)
        s, b, osc, cam, ctx, rank = setup(sc, W, W)
        print("  bvh", b.info(), flush=True)
        rays_o = osc.camera_rays(cam)
        rays_g = ctx.camera_rays()
        print("  camera rays bit-equal:", rays_o.tobytes() == rays_g.tobytes(), flush=True)
        oi, ot = osc.primary_hits(cam)
        for mode in (api.TRAVERSAL_REF_BVH2, api.TRAVERSAL_WIDE):
            ctx.set_options(traversal=mode, max_depth=5)
            t0 = time.time(); ids, ts = ctx.primary_hits(); dt = time.time() - t0
            bad = cmp_hits(f"primary mode={mode} ({dt*1e3:.1f} ms incl. copies)", ids, ts, oi, ot)
            if len(bad) and mode == api.TRAVERSAL_WIDE:
                bi, bt = osc.brute_force(rays_o[bad[:64]], rank)
                k = bad[:64]
                print("   wide==brute on", int(((bi == ids[k]) & (bt.view(np.uint32) == ts[k].view(np.uint32))).sum()), "of", len(k))
        # incoherent rays
        rng = np.random.default_rng(3)
        vv = s.arrays()[0].view(np.float32).reshape(-1, 3); lo, hi = vv.min(0), vv.max(0)
        n = 200000
        o = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32); d = rng.normal(size=(n, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        rr = np.concatenate([o, d], axis=1).astype(np.float32)
        oi2, ot2 = osc.trace_rays(rr)
        for mode in (api.TRAVERSAL_REF_BVH2, api.TRAVERSAL_WIDE):
            ctx.set_options(traversal=mode)
            ids, ts = ctx.trace_rays(rr)
            cmp_hits(f"random rays mode={mode}", ids, ts, oi2, ot2)
        # radiance vs oracle, small
        Ws = 128 if name == "atrium" else 256
        cam_s = api.make_camera(Ws, Ws, **sc.camera)
        ctx.set_camera(cam_s)
        spp = 16
        t0 = time.time(); om, osum, orays = osc.render_pt(cam_s, 5, 1234, 0, spp); dto = time.time() - t0
        for mode in (api.TRAVERSAL_WIDE, api.TRAVERSAL_REF_BVH2):
            ctx.set_options(traversal=mode, max_depth=5, seed=1234, collect_stats=1)
            ctx.reset_accum(); ctx.reset_stats()
            ctx.render(0, spp); img = ctx.read_accum(True)
            st = ctx.stats()
            ref = (osum / spp)
            diff = img.astype(np.float64) - ref
            rmse = np.sqrt((diff ** 2).mean()); rel = rmse / max(ref.mean(), 1e-12)
            exact = np.mean(np.all(np.abs(diff) <= 1e-4 * (1 + np.abs(ref)), axis=1))
            print(f"  radiance mode={mode}: mean gpu {img.mean():.6f} oracle {ref.mean():.6f} rmse {rmse:.5f} rel {rel:.5f} px-close {exact:.4f} rays gpu {st['rays']} oracle {orays} (oracle {dto:.2f}s)", flush=True)
        # timing at full config
        if name == "atrium":
            ctx.set_camera(cam)
            for mode, spp_t in ((api.TRAVERSAL_WIDE, 64), (api.TRAVERSAL_REF_BVH2, 2)):
                ctx.set_options(traversal=mode, max_depth=5, collect_stats=0, time_stages=1)
                ctx.reset_accum(); ctx.render(0, 2); ctx.sync(); ctx.reset_stats(); ctx.reset_accum()
                t0 = time.time(); ctx.render(0, spp_t); ctx.sync(); dt = time.time() - t0
                st = ctx.stats()
                print(f"  TIMING mode={mode} {W}x{W} {spp_t}spp: {dt*1e3:.1f} ms wall, rays {st['rays']} -> {st['rays']/dt/1e6:.1f} Mrays/s, {W*W*spp_t/dt/1e6:.1f} Msamples/s; stages ms raygen {st['ms_raygen']:.1f} extend {st['ms_extend']:.1f} shade {st['ms_shade']:.1f} accum {st['ms_accum']:.1f} launches {st['kernel_launches']}", flush=True)
            ctx.set_options(traversal=api.TRAVERSAL_WIDE, collect_stats=1, time_stages=0)
            ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 4); st = ctx.stats()
            print(f"  STATS wide: nodes/ray {st['nodes_visited']/st['rays']:.2f} tris/ray {st['tris_tested']/st['rays']:.2f} rays/sample {st['rays']/st['samples']:.3f}")
            img = ctx.read_accum(True)
            api.write_ppm(os.path.join(ROOT, "gpurun_out", "atrium_4spp.ppm"), img, W, W)
        ctx.close()

if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    main()
