"""TEST INFRASTRUCTURE: ctypes binding of oracle/_ref/libcupt_ref.so — the real reference
(CulDeVu/CUDAPathTracer, patched per oracle/build_ref.sh) running on host cores.  Only tests/, bench.py's
reference arm and __graft_entry__.smoke() may import this."""
from __future__ import annotations

import ctypes as C
import os
import tempfile

import numpy as np

from cudapathtracer_b200.api import BVH2_NODE, CAMERA, MATERIAL, TRIANGLE, VEC3

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libcupt_ref.so")


def available() -> bool:
    return os.path.exists(REF_SO)


_lib = None


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(REF_SO)
        l.ref_total_light_area.restype = C.c_float
        l.ref_scene_reset.restype = None
        _lib = l
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def struct_sizes():
    out = (C.c_int * 10)()
    lib().ref_struct_sizes(out)
    names = ["vec3", "triangle", "materialDesc", "sceneDesc", "camera", "BVH_array_node", "BVH_array", "color", "ray", "AABB"]
    return dict(zip(names, out))


def reset():
    lib().ref_scene_reset()


def load_obj_text(obj_text: str, mtl_text: str, origin=(0.0, 0.0, 0.0), scale=1.0, flip=False, mtl_name="scene.mtl"):
    """write the texts under <tmp>/models/ and call the reference's loadOBJ from <tmp> (MTL base path is the
    hard-coded "models/", modelLoader.h:132)"""
    with tempfile.TemporaryDirectory() as d:
        os.makedirs(os.path.join(d, "models"))
        with open(os.path.join(d, "models", "scene.obj"), "w") as f:
            f.write(obj_text)
        with open(os.path.join(d, "models", mtl_name), "w") as f:
            f.write(mtl_text)
        rc = lib().ref_load_obj(d.encode(), b"models/scene.obj", C.c_float(origin[0]), C.c_float(origin[1]),
                                C.c_float(origin[2]), C.c_float(scale), int(flip))
        assert rc == 0, rc


def set_arrays(verts, tris, mats, lights, area):
    verts = np.ascontiguousarray(verts, dtype=VEC3)
    tris = np.ascontiguousarray(tris, dtype=TRIANGLE)
    mats = np.ascontiguousarray(mats, dtype=MATERIAL)
    lights = np.ascontiguousarray(lights, dtype=np.int32)
    lib().ref_scene_set_arrays(_p(verts), len(verts), _p(tris), len(tris), _p(mats), len(mats), _p(lights), len(lights),
                               C.c_float(area))


def arrays():
    c = (C.c_int * 4)()
    lib().ref_scene_counts(c)
    verts, tris = np.zeros(c[0], dtype=VEC3), np.zeros(c[1], dtype=TRIANGLE)
    mats, lights = np.zeros(c[2], dtype=MATERIAL), np.zeros(c[3], dtype=np.int32)
    lib().ref_scene_get_arrays(_p(verts), _p(tris), _p(mats), _p(lights))
    return verts, tris, mats, lights, np.float32(lib().ref_total_light_area())


def build_bvh():
    """the reference's buildBVH(): (nodes, depth).  stdout chatter of the reference is left alone."""
    lib().ref_build_bvh()
    size, depth = C.c_int(), C.c_int()
    lib().ref_bvh_info(C.byref(size), C.byref(depth))
    nodes = np.zeros(size.value, dtype=BVH2_NODE)
    lib().ref_bvh_get(_p(nodes))
    return nodes, depth.value


def camera_rays(cam: np.ndarray, u1=0.5, u2=0.5):
    cam = np.ascontiguousarray(cam, dtype=CAMERA)
    w, h = int(cam["pxlWidth"][0]), int(cam["pxlHeight"][0])
    rays = np.zeros((h * w, 6), dtype=np.float32)
    lib().ref_camera_rays(_p(cam), C.c_float(u1), C.c_float(u2), _p(rays))
    return rays


def trace_rays(rays: np.ndarray, nthreads=8):
    rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
    ids = np.zeros(len(rays), dtype=np.int32)
    ts = np.zeros(len(rays), dtype=np.float32)
    lib().ref_trace_rays(_p(rays), C.c_longlong(len(rays)), nthreads, _p(ids), _p(ts))
    return ids, ts


def primary_hits(cam: np.ndarray, u1=0.5, u2=0.5, nthreads=8):
    cam = np.ascontiguousarray(cam, dtype=CAMERA)
    w, h = int(cam["pxlWidth"][0]), int(cam["pxlHeight"][0])
    ids = np.zeros(h * w, dtype=np.int32)
    ts = np.zeros(h * w, dtype=np.float32)
    lib().ref_primary_hits(_p(cam), C.c_float(u1), C.c_float(u2), nthreads, _p(ids), _p(ts))
    return ids, ts


def render(cam: np.ndarray, integrator=2, num_bounces=5, rng=1, seed=1234, first_sample=0, num_samples=1, nthreads=8):
    """returns (mean[P,3] f64 — the reference's running mean, sum[P,3] f64, rays counted)"""
    cam = np.ascontiguousarray(cam, dtype=CAMERA)
    w, h = int(cam["pxlWidth"][0]), int(cam["pxlHeight"][0])
    mean = np.zeros((h * w, 3), dtype=np.float64)
    ssum = np.zeros((h * w, 3), dtype=np.float64)
    rays = C.c_ulonglong()
    rc = lib().ref_render(_p(cam), integrator, num_bounces, rng, C.c_ulonglong(seed), first_sample, num_samples, nthreads,
                          _p(mean), _p(ssum), C.byref(rays))
    assert rc == 0, rc
    return mean, ssum, rays.value


def philox_uniforms(seed, x, y, sample, n):
    out = np.zeros(n, dtype=np.float32)
    lib().ref_philox_uniforms(C.c_ulonglong(seed), x, y, sample, n, _p(out))
    return out


def tonemap(rgb: np.ndarray):
    rgb = np.ascontiguousarray(rgb, dtype=np.float64).reshape(-1, 3)
    out = np.zeros((len(rgb), 3), dtype=np.int32)
    lib().ref_tonemap(_p(rgb), C.c_longlong(len(rgb)), _p(out))
    return out
