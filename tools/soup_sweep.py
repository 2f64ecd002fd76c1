"""Config 4 (BASELINE.json): triangle-soup BVH stress sweep, incoherent diffuse secondary rays, 1024^2, 16 spp, depth 5.
GPU box tool; writes one JSON line per size to gpurun_out/config4_soup.jsonl.  Sizes from SIZES (millions, comma list)."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cudapathtracer_b200 import api, scenes
import oraclebind

PEAK = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
W, SPP, DEPTH = 1024, 16, 5
BVH = os.environ.get("BVH", "device")
RADIUS = int(os.environ.get("RADIUS", "0"))


def main():
    sizes = [float(x) for x in os.environ.get("SIZES", "1,3,10").split(",")]
    out = open(os.path.join(ROOT, "gpurun_out", "config4_soup.jsonl"), "a")
    ctx = api.Context(0)
    for m in sizes:
        n = int(m * 1e6)
        t0 = time.time(); sc = scenes.triangle_soup(n); tg = time.time() - t0
        s = api.SceneHandle(); s.add_scene(sc)
        del sc.shapes[:]
        cam = api.make_camera(W, W, **sc.camera)
        ctx.set_scene(s); ctx.set_camera(cam)
        b = None
        if BVH == "device":   # built on the GPU from the uploaded scene (csrc/bvh_device.cu)
            ctx.build_bvh_device(radius=RADIUS)   # warm-up (first-use allocations), then the timed build
            t0 = time.time(); info = ctx.build_bvh_device(radius=RADIUS); tb = time.time() - t0
        else:
            t0 = time.time(); b = s.build_bvh(); tb = time.time() - t0
            info = b.info()
            ctx.set_bvh(b)
        ctx.set_options(max_depth=DEPTH, seed=1234, traversal=api.TRAVERSAL_WIDE, collect_stats=1, time_stages=0)
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 2); st = ctx.stats()
        npr, tpr = st["nodes_visited"] / st["rays"], st["tris_tested"] / st["rays"]
        ctx.set_options(collect_stats=0, time_stages=1)
        ctx.reset_accum(); ctx.render(0, 2); ctx.sync()
        best, stb = 1e9, None
        for _ in range(3):
            ctx.reset_accum(); ctx.reset_stats()
            t0 = time.perf_counter(); ctx.render(0, SPP); ctx.sync(); dt = time.perf_counter() - t0
            if dt < best: best, stb = dt, ctx.stats()
        rays = stb["rays"]
        bpr = npr * 80 + tpr * 48 + 44
        ext_gbs = rays * bpr / (stb["ms_extend"] * 1e-3) / 1e9
        rec = {"config": 4, "triangles": n, "gen_s": round(tg, 2), "bvh": BVH, "ploc_radius": RADIUS, "bvh_build_s": round(tb, 3), "bvh_build_ms_device": {k: round(v, 2) for k, v in info.items() if k.startswith("ms_")}, "nodes": info["num_nodes"], "depth": info["depth"],
               "bvh_mb": round((info["num_nodes"] * 80 + info["num_tri_records"] * 48) / 1e6, 1), "image": f"{W}x{W}", "spp": SPP, "max_depth": DEPTH,
               "ms": round(best * 1e3, 2), "mrays_per_s": round(rays / best / 1e6, 1), "rays_per_sample": round(rays / (W * W * SPP), 3),
               "nodes_per_ray": round(npr, 2), "tris_per_ray": round(tpr, 2), "bytes_per_ray": round(bpr, 1), "ms_extend": round(stb["ms_extend"], 2),
               "extend_algorithmic_gbs": round(ext_gbs, 1), "frac_of_measured_hbm_peak": round(ext_gbs / PEAK, 3)}
        # parity property at full size: a sample of rays against the brute force over ALL triangles (oracle), bit for bit
        nb = int(os.environ.get("BRUTE_RAYS", "400")) if n <= 10_000_000 else 0
        if nb:
            v, t, mm, l = s.arrays()
            rng = np.random.default_rng(3)
            o = rng.uniform(-50, 50, size=(nb, 3)); d = rng.normal(size=(nb, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
            rr = np.concatenate([o, d], axis=1).astype(np.float32)
            osc = oraclebind.OracleScene(v, t, mm, l, s.total_light_area, np.zeros(1, api.BVH2_NODE))
            bi, bt = osc.brute_force(rr, None)
            gi, gt = ctx.trace_rays(rr)
            rec["brute_force_rays"] = nb
            rec["brute_force_mismatches"] = int(((bi != gi) | (bt.view(np.uint32) != gt.view(np.uint32))).sum())
            rec["brute_force_hits"] = int((bi >= 0).sum())
        print(json.dumps(rec), flush=True)
        out.write(json.dumps(rec) + "\n"); out.flush()
        if b is not None: b.close()
        s.close()
    ctx.close()


if __name__ == "__main__":
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    main()
