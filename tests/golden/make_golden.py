"""Generates tests/golden/*.npz from the REAL reference (oracle/_ref/libcupt_ref.so, built from /root/reference by
oracle/build_ref.sh).  Run here (needs /root/reference); the vectors travel, the reference does not.

For each small case: the reference's own loadOBJ output (via OBJ text), buildBVH() node array, primary hit ids and
t bits of cameraRay + trace, and the f64 radiance sums of radianceAlongSingleStep2 and of the live
radianceAlongSingleStep (BDPT) with the keyed Philox stream.
    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import refbind  # noqa: E402
from conftest import CASES, build_case  # noqa: E402
from cudapathtracer_b200 import api, scenes  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    assert refbind.available(), "build oracle/_ref first (oracle/build_ref.sh)"
    for name in CASES:
        sc, w, h = build_case(name)
        obj, mtl = scenes.to_obj(sc)
        refbind.reset()
        refbind.load_obj_text(obj, mtl)
        verts, tris, mats, lights, area = refbind.arrays()
        nodes, depth = refbind.build_bvh()
        cam = api.make_camera(w, h, **sc.camera)
        ids, ts = refbind.primary_hits(cam, 0.5, 0.5)
        rays = refbind.camera_rays(cam, 0.5, 0.5)
        out = dict(verts=verts, tris=tris, mats=mats, lights=lights, area=np.float32(area), nodes=nodes, depth=np.int32(depth),
                   primary_ids=ids, primary_t_bits=ts.view(np.uint32), camera_rays_sha=np.array(sha(rays)))
        for nb, spp, seed in ((5, 4, 1234), (3, 2, 99)):
            mean, ssum, nrays = refbind.render(cam, integrator=2, num_bounces=nb, rng=1, seed=seed, first_sample=0, num_samples=spp)
            out[f"pt_sum_nb{nb}_spp{spp}_seed{seed}"] = ssum
            out[f"pt_mean_nb{nb}_spp{spp}_seed{seed}"] = mean
            out[f"pt_rays_nb{nb}_spp{spp}_seed{seed}"] = np.uint64(nrays)
        # the LIVE integrator radianceAlongSingleStep (fixed-length BDPT, kernel.cu:217-415), same keyed Philox stream
        mean, ssum, nrays = refbind.render(cam, integrator=1, num_bounces=3, rng=1, seed=1234, first_sample=0, num_samples=4)
        out["bdpt_sum_spp4_seed1234"], out["bdpt_mean_spp4_seed1234"], out["bdpt_rays_spp4_seed1234"] = ssum, mean, np.uint64(nrays)
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
        print(name, "tris", len(tris), "nodes", len(nodes), "depth", depth, "hits", int((ids >= 0).sum()), "/", len(ids))
    ph = {f"{s}_{x}_{y}_{k}": refbind.philox_uniforms(s, x, y, k, 19) for s, x, y, k in ((1234, 0, 0, 0), (1234, 5, 7, 3), (2**40 + 17, 1023, 2159, 1023))}
    x = np.abs(np.random.default_rng(0).normal(size=(256, 3))) * 3
    np.savez_compressed(os.path.join(HERE, "misc.npz"), tonemap_in=x, tonemap_out=refbind.tonemap(x), **{"philox_" + k: v for k, v in ph.items()})
    print("sizes", refbind.struct_sizes())


if __name__ == "__main__":
    main()
