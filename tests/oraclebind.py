"""TEST INFRASTRUCTURE: ctypes binding of oracle/liboracle.so (the C restatement, oracle/oracle.c).
Only tests/, bench.py's cpu_baseline / reference arm and __graft_entry__.smoke() may import this."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from cudapathtracer_b200.api import BVH2_NODE, CAMERA, MATERIAL, SPHERE, TRIANGLE, VEC3

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "liboracle.so")


class _Scene(C.Structure):
    _fields_ = [("verts", C.c_void_p), ("tris", C.c_void_p), ("mats", C.c_void_p), ("lights", C.c_void_p),
                ("numVerts", C.c_int32), ("numTris", C.c_int32), ("numMats", C.c_int32), ("numLights", C.c_int32),
                ("totalLightArea", C.c_float), ("nodes", C.c_void_p), ("numNodes", C.c_int32),
                ("spheres", C.c_void_p), ("numSpheres", C.c_int32), ("mat_specular", C.c_void_p)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(ORACLE_SO)
        l.orc_render_pt.restype = C.c_ulonglong
        l.orc_render_bdpt.restype = C.c_ulonglong
        _lib = l
    return _lib


def available() -> bool:
    return os.path.exists(ORACLE_SO)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleScene:
    """scene arrays in reference layout + a reference-layout BVH2 node array"""

    def __init__(self, verts, tris, mats, lights, total_light_area, nodes, spheres=None, specular=None):
        self.verts = np.ascontiguousarray(verts, dtype=VEC3)
        self.tris = np.ascontiguousarray(tris, dtype=TRIANGLE)
        self.mats = np.ascontiguousarray(mats, dtype=MATERIAL)
        self.lights = np.ascontiguousarray(lights, dtype=np.int32)
        self.nodes = np.ascontiguousarray(nodes, dtype=BVH2_NODE)
        self.c = _Scene(self.verts.ctypes.data, self.tris.ctypes.data, self.mats.ctypes.data, self.lights.ctypes.data,
                        len(self.verts), len(self.tris), len(self.mats), len(self.lights),
                        float(total_light_area), self.nodes.ctypes.data, len(self.nodes), None, 0, None)
        if specular is not None and len(specular):   # material indices that are perfect mirrors (extension, unpinned)
            self.specular = np.zeros(len(self.mats), dtype=np.uint8)
            self.specular[np.asarray(specular, dtype=np.int64)] = 1
            self.c.mat_specular = self.specular.ctypes.data
        if spheres is not None and len(spheres):
            self.spheres = np.ascontiguousarray(spheres, dtype=SPHERE)
            self.c.spheres, self.c.numSpheres = self.spheres.ctypes.data, len(self.spheres)

    def camera_rays(self, cam, u1=0.5, u2=0.5):
        cam = np.ascontiguousarray(cam, dtype=CAMERA)
        n = int(cam["pxlWidth"][0]) * int(cam["pxlHeight"][0])
        rays = np.zeros((n, 6), dtype=np.float32)
        lib().orc_camera_rays(_p(cam), C.c_float(u1), C.c_float(u2), _p(rays))
        return rays

    def primary_hits(self, cam, u1=0.5, u2=0.5):
        cam = np.ascontiguousarray(cam, dtype=CAMERA)
        n = int(cam["pxlWidth"][0]) * int(cam["pxlHeight"][0])
        ids, ts = np.zeros(n, dtype=np.int32), np.zeros(n, dtype=np.float32)
        lib().orc_primary_hits(C.byref(self.c), _p(cam), C.c_float(u1), C.c_float(u2), _p(ids), _p(ts))
        return ids, ts

    def trace_rays(self, rays):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        ids, ts = np.zeros(len(rays), dtype=np.int32), np.zeros(len(rays), dtype=np.float32)
        lib().orc_trace_rays(C.byref(self.c), _p(rays), C.c_longlong(len(rays)), _p(ids), _p(ts))
        return ids, ts

    def brute_force(self, rays, rank=None):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        if rank is not None:
            rank = np.ascontiguousarray(rank, dtype=np.uint32)
        ids, ts = np.zeros(len(rays), dtype=np.int32), np.zeros(len(rays), dtype=np.float32)
        lib().orc_brute_force(C.byref(self.c), _p(rank), _p(rays), C.c_longlong(len(rays)), _p(ids), _p(ts))
        return ids, ts

    def hits_outside_box(self, rays, rel_tol=4.8e-7):
        """per ray: does any triangle report a hit outside its own bounding box as the builders pad it (4 ulp per axis) — rounding
        noise of a (nearly) degenerate triangle, which a tree may or may not see?"""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        flags = np.zeros(len(rays), dtype=np.uint8)
        lib().orc_hits_outside_box(C.byref(self.c), _p(rays), C.c_longlong(len(rays)), C.c_double(rel_tol), _p(flags))
        return flags.astype(bool)

    def render_pt(self, cam, num_bounces=5, seed=1234, first_sample=0, num_samples=1):
        """(mean[P,3] f64, sum[P,3] f64, rays)"""
        cam = np.ascontiguousarray(cam, dtype=CAMERA)
        n = int(cam["pxlWidth"][0]) * int(cam["pxlHeight"][0])
        mean, ssum = np.zeros((n, 3)), np.zeros((n, 3))
        rays = lib().orc_render_pt(C.byref(self.c), _p(cam), num_bounces, C.c_uint64(seed), first_sample, num_samples,
                                   _p(mean), _p(ssum))
        return mean, ssum, int(rays)


def _render_bdpt(self, cam, seed=1234, first_sample=0, num_samples=1):
    """live integrator radianceAlongSingleStep (kernel.cu:217-415): (mean, sum, rays)"""
    cam = np.ascontiguousarray(cam, dtype=CAMERA)
    n = int(cam["pxlWidth"][0]) * int(cam["pxlHeight"][0])
    mean, ssum = np.zeros((n, 3)), np.zeros((n, 3))
    rays = lib().orc_render_bdpt(C.byref(self.c), _p(cam), C.c_uint64(seed), first_sample, num_samples, _p(mean), _p(ssum))
    return mean, ssum, int(rays)


OracleScene.render_bdpt = _render_bdpt


def philox_uniforms(seed, x, y, sample, n):
    out = np.zeros(n, dtype=np.float32)
    lib().orc_philox_uniforms(C.c_uint64(seed), x, y, sample, n, _p(out))
    return out


def tonemap(rgb):
    rgb = np.ascontiguousarray(rgb, dtype=np.float64).reshape(-1, 3)
    out = np.zeros((len(rgb), 3), dtype=np.int32)
    lib().orc_tonemap(_p(rgb), C.c_longlong(len(rgb)), _p(out))
    return out
