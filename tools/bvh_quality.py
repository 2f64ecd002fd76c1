"""Developer tool (no GPU): node visits / primitive tests per ray of a wide BVH on config 2, through the host emulation of
the traversal header (tests/emul), on a fixed mix of primary and cosine-bounce rays.  Lets builder changes be judged
before a GPU run:  python tools/bvh_quality.py [key=value ...]   (keys: build_bvh's keyword arguments)"""
import ctypes as C, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cudapathtracer_b200 import api, scenes

EMUL = C.CDLL(os.path.join(ROOT, "tests", "emul", "libemul.so"))
P = lambda a: a.ctypes.data_as(C.c_void_p)

def trace(nodes, prims, depth, rays):
    ids = np.zeros(len(rays), np.int32); ts = np.zeros(len(rays), np.float32); cnt = (C.c_ulonglong * 2)()
    EMUL.emul_trace_wide_fast(P(nodes), P(prims), depth, P(rays), C.c_longlong(len(rays)), P(ids), P(ts), cnt)
    return ids, ts, cnt[0], cnt[1]

def ray_mix(s, sc, nodes, prims, depth, res=192, bounces=2, seed=3):
    """primary rays of a res x res image and `bounces` generations of cosine-weighted bounce rays off the hit points"""
    import oraclebind
    v, t, m, l = s.arrays()
    cam = api.make_camera(res, res, **sc.camera)
    osc = oraclebind.OracleScene(v, t, m, l, s.total_light_area, np.zeros(1, api.BVH2_NODE))
    rays = osc.camera_rays(cam)
    rng = np.random.default_rng(seed)
    out = [rays]
    for _ in range(bounces):
        ids, ts, _, _ = trace(nodes, prims, depth, rays)
        ok = ids >= 0
        pos = rays[ok, :3] + rays[ok, 3:] * (ts[ok, None] - 1e-3)
        n = t["norm"][ids[ok]].astype(np.float64)
        n /= np.linalg.norm(n, axis=1, keepdims=True)
        n[np.einsum("ij,ij->i", n, rays[ok, 3:]) > 0] *= -1
        u1, u2 = rng.random(len(n)), rng.random(len(n))
        a = np.where(np.abs(n[:, :1]) > 0.9, [[0, 1, 0]], [[1, 0, 0]])
        tx = np.cross(n, a); tx /= np.linalg.norm(tx, axis=1, keepdims=True); ty = np.cross(n, tx)
        r, th = np.sqrt(u1), 2 * np.pi * u2
        d = n * np.sqrt(1 - u1)[:, None] + tx * (r * np.cos(th))[:, None] + ty * (r * np.sin(th))[:, None]
        rays = np.concatenate([pos, d / np.linalg.norm(d, axis=1, keepdims=True)], 1).astype(np.float32)
        out.append(rays)
    return out

def main():
    kw = {k: int(v) for k, v in (a.split("=") for a in sys.argv[1:])}
    sc = scenes.atrium()
    s = api.SceneHandle(); s.add_scene(sc)
    base = s.build_bvh()
    bn, bp = base.arrays()
    mix = ray_mix(s, sc, bn, bp, base.info()["depth"])
    t0 = time.perf_counter(); b = s.build_bvh(**kw); dt = time.perf_counter() - t0
    nodes, prims = b.arrays(); info = b.info()
    line = f"{kw}: build {dt:.2f} s, {info['num_nodes']} nodes, {info['num_tri_records']} records, depth {info['depth']}, SAH {info['sah_cost']:.2f};"
    tn = tt = tr = 0
    want = None
    for g, rays in enumerate(mix):
        ids, ts, nn, nt = trace(nodes, prims, info["depth"], rays)
        bi, bt, _, _ = trace(bn, bp, base.info()["depth"], rays)
        assert np.array_equal(ids, bi) and ts.tobytes() == bt.tobytes(), "hits differ from the default tree's"
        line += f"  gen{g} nodes/ray {nn / len(rays):.2f} prims/ray {nt / len(rays):.2f}"
        tn += nn; tt += nt; tr += len(rays)
    allrays = np.concatenate(mix)
    cnt = (C.c_ulonglong * 5)()
    EMUL.emul_trace_wide_count_empty(P(nodes), P(prims), info["depth"], P(allrays), C.c_longlong(len(allrays)), cnt)
    print(line + f"  ALL nodes/ray {tn / tr:.3f} prims/ray {tt / tr:.3f} empty node steps/ray {cnt[2] / len(allrays):.3f}; "
          f"steps in nodes containing the ray origin {cnt[3] / len(allrays):.3f} (empty ones {cnt[4] / len(allrays):.3f})", flush=True)

if __name__ == "__main__":
    main()
