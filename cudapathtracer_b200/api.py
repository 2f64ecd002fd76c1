"""ctypes binding of libcupt.so (include/cupt.h) — the Python face of the C-ABI used by tests and bench.

There is no CPU fallback: everything that renders or traces goes through the CUDA kernels in libcupt.so
and raises CuptError when the library or a CUDA device is missing.  Host-side entry points (scene ingest,
the reference-layout BVH2, the wide-BVH builder, PPM output) work without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CUPT_LIB") or os.path.join(_HERE, "libcupt.so")   # CUPT_LIB: an experiment build of the same library (tools/)

# numpy views of the reference PODs (sizes pinned in tests/test_abi.py)
VEC3 = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4")])
TRIANGLE = np.dtype([("v0", "<i4"), ("v1", "<i4"), ("v2", "<i4"), ("norm", "<f4", (3,)), ("mat", "<i4")])
MATERIAL = np.dtype([("albedo", "<f8", (3,)), ("emission", "<f8", (3,))])
CAMERA = np.dtype([("pos", "<f4", (3,)), ("distFromFilm", "<f4"), ("focalLength", "<f4"), ("radius", "<f4"),
                   ("pxlWidth", "<i4"), ("pxlHeight", "<i4")])
BVH2_NODE = np.dtype([("lo", "<f4", (3,)), ("hi", "<f4", (3,)), ("left", "<u4"), ("right", "<u4")])
SPHERE = np.dtype([("pos", "<f4", (3,)), ("rad", "<f4"), ("diffuse", "<f8", (3,)), ("emm", "<f8", (3,))])
assert (VEC3.itemsize, TRIANGLE.itemsize, MATERIAL.itemsize, CAMERA.itemsize, BVH2_NODE.itemsize, SPHERE.itemsize) == (12, 28, 48, 32, 32, 64)

LEAF_FLAG = 0x80000000
MAX_FLOAT = np.float32(100000.0)
INTEGRATOR_PT, INTEGRATOR_BDPT = 2, 1
TRAVERSAL_WIDE, TRAVERSAL_REF_BVH2, TRAVERSAL_WIDE_STRICT = 0, 1, 2
TRI_REFEXACT, TRI_WATERTIGHT = 0, 1
OBJ_FACE_MATERIALS, OBJ_DEDUP_MATERIALS = 1, 2
NOALLOC_MAX_RAYS = 65536   # CUPT_NOALLOC_MAX_RAYS


class CuptError(RuntimeError):
    pass


class BvhOpts(C.Structure):
    _fields_ = [("sah_bins", C.c_int), ("max_leaf_tris", C.c_int), ("threads", C.c_int), ("tie_rank", C.c_void_p), ("ploc_radius", C.c_int), ("reinsert_passes", C.c_int), ("greedy_collapse", C.c_int)]


class BvhInfo(C.Structure):
    _fields_ = [("num_nodes", C.c_int64), ("num_tri_records", C.c_int64), ("num_tris", C.c_int64),
                ("num_spheres", C.c_int64), ("depth", C.c_int32), ("node_bytes", C.c_int32), ("tri_bytes", C.c_int32),
                ("sah_cost", C.c_double), ("build_seconds", C.c_double)]


class RenderOpts(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("integrator", C.c_int32), ("traversal", C.c_int32), ("tri_test", C.c_int32),
                ("seed", C.c_uint64), ("samples_per_pass", C.c_int32), ("collect_stats", C.c_int32),
                ("time_stages", C.c_int32), ("fetch_threshold", C.c_int32), ("defer_pairs", C.c_int32),
                ("share_primary_hits", C.c_int32), ("no_terminal_stop", C.c_int32), ("record_retraces", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("rays", C.c_uint64), ("samples", C.c_uint64), ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("ms_raygen", C.c_double), ("ms_extend", C.c_double),
                ("ms_shade", C.c_double), ("ms_accum", C.c_double), ("extend_launches", C.c_uint64), ("strict_retraces", C.c_uint64)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_lib = None


def load_library(path: Optional[str] = None):
    """dlopen libcupt.so (built in-tree by __graft_entry__.build()); fails loudly when it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise CuptError(f"{p} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                        "(there is no CPU or PyTorch fallback for the hot path)")
    lib = C.CDLL(p, mode=C.RTLD_GLOBAL)
    lib.cupt_last_error.restype = C.c_char_p
    lib.cupt_scene_total_light_area.restype = C.c_float
    lib.cupt_scene_total_light_area.argtypes = [C.c_void_p]
    for name in ("cupt_scene_destroy", "cupt_bvh_destroy", "cupt_destroy", "cupt_bvh2_free"):
        getattr(lib, name).restype = None
        getattr(lib, name).argtypes = [C.c_void_p]
    _lib = lib
    return lib


def _check(rc: int):
    if rc != 0:
        raise CuptError(f"cupt error {rc}: {load_library().cupt_last_error().decode(errors='replace')}")


def _ptr(a: Optional[np.ndarray]):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_camera(width: int, height: int, pos=(0.0, 1.0, 3.0), distFromFilm=1.0, focalLength=3.0, radius=0.0) -> np.ndarray:
    cam = np.zeros(1, dtype=CAMERA)
    cam["pos"][0] = pos
    cam["distFromFilm"], cam["focalLength"], cam["radius"] = distFromFilm, focalLength, radius
    cam["pxlWidth"], cam["pxlHeight"] = width, height
    return cam


def spheres_array(spheres) -> np.ndarray:
    """[(centre, radius, diffuse, emm)] -> SPHERE records; colours pass through float like MTL values do"""
    out = np.zeros(len(spheres), dtype=SPHERE)
    for i, (c, r, kd, ke) in enumerate(spheres):
        out["pos"][i] = c
        out["rad"][i] = r
        out["diffuse"][i] = np.asarray(kd, dtype=np.float32)
        out["emm"][i] = np.asarray(ke, dtype=np.float32)
    return out


class SceneHandle:
    """cupt_scene: the reference's globals verts/tris/mats/lights/totalLightArea (modelLoader.h:43-47)."""

    def __init__(self):
        self.h = None
        self.lib = load_library()
        h = C.c_void_p()
        _check(self.lib.cupt_scene_create(C.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            self.lib.cupt_scene_destroy(self.h)
            self.h = None

    __del__ = close

    def clear(self):
        _check(self.lib.cupt_scene_clear(self.h))

    def load_obj(self, path: str, mtl_dir: Optional[str] = None, origin=(0.0, 0.0, 0.0), scale: float = 1.0, flip: bool = False,
                 face_materials: bool = False, dedup_materials: bool = False):
        """loadOBJ(filename, origin, scale, flipNormals) — modelLoader.h:125; the two flags go beyond the reference
        (per-face materials, de-duplicated material table: SURVEY 8f rank 2)"""
        o = (C.c_float * 3)(*origin)
        flags = (OBJ_FACE_MATERIALS if face_materials else 0) | (OBJ_DEDUP_MATERIALS if dedup_materials else 0)
        _check(self.lib.cupt_scene_load_obj_ex(self.h, path.encode(), None if mtl_dir is None else mtl_dir.encode(),
                                               o, C.c_float(scale), int(flip), flags))

    def save(self, path: str):
        """binary scene cache (cupt_scene_save)"""
        _check(self.lib.cupt_scene_save(self.h, path.encode()))

    def load(self, path: str):
        """replace the contents with a scene cache written by save()"""
        _check(self.lib.cupt_scene_load(self.h, path.encode()))

    def add_materials(self, mats: np.ndarray, duplicate_table: bool = True) -> int:
        mats = np.ascontiguousarray(mats, dtype=MATERIAL)
        first = C.c_int()
        _check(self.lib.cupt_scene_add_materials(self.h, _ptr(mats), len(mats), int(duplicate_table), C.byref(first)))
        return first.value

    def add_shape(self, positions, indices, material: int, origin=(0.0, 0.0, 0.0), scale=1.0, flip=False):
        positions = np.ascontiguousarray(positions, dtype=np.float32).reshape(-1, 3)
        indices = np.ascontiguousarray(indices, dtype=np.uint32).reshape(-1, 3)
        o = (C.c_float * 3)(*origin)
        _check(self.lib.cupt_scene_add_shape(self.h, _ptr(positions), len(positions), _ptr(indices), len(indices),
                                             int(material), o, C.c_float(scale), int(flip)))

    def add_scene(self, scene, origin=(0.0, 0.0, 0.0), scale=1.0, flip=False, duplicate_table=True):
        """append a cudapathtracer_b200.scenes.Scene the way loadOBJ appends an OBJ file"""
        mats = np.zeros(len(scene.materials), dtype=MATERIAL)
        for i, (_, kd, ke) in enumerate(scene.materials):
            mats["albedo"][i] = np.asarray(kd, dtype=np.float32)   # MTL values pass through float (tinyobj material_t)
            mats["emission"][i] = np.asarray(ke, dtype=np.float32)
        first = self.add_materials(mats, duplicate_table)
        for sh in scene.shapes:
            self.add_shape(sh.positions, sh.indices, sh.material + first, origin, scale, flip)
        if getattr(scene, "spheres", None):
            self.add_spheres(spheres_array(scene.spheres))
        for m in getattr(scene, "specular", ()):
            self.set_specular(m + first)

    def set_specular(self, material: int, specular: bool = True):
        """mark a material as a perfect mirror (extension for BASELINE config 5; the reference is diffuse only)"""
        _check(self.lib.cupt_scene_set_specular(self.h, int(material), int(specular)))

    def add_spheres(self, spheres: np.ndarray):
        spheres = np.ascontiguousarray(spheres, dtype=SPHERE)
        _check(self.lib.cupt_scene_add_spheres(self.h, _ptr(spheres), len(spheres)))

    def set_arrays(self, verts, tris, mats, lights, total_light_area: float):
        verts = np.ascontiguousarray(verts, dtype=VEC3)
        tris = np.ascontiguousarray(tris, dtype=TRIANGLE)
        mats = np.ascontiguousarray(mats, dtype=MATERIAL)
        lights = np.ascontiguousarray(lights, dtype=np.int32)
        _check(self.lib.cupt_scene_set_arrays(self.h, _ptr(verts), len(verts), _ptr(tris), len(tris), _ptr(mats), len(mats),
                                              _ptr(lights), len(lights), C.c_float(total_light_area)))

    def counts(self):
        out = (C.c_int * 5)()
        _check(self.lib.cupt_scene_counts(self.h, out))
        return tuple(out)

    @property
    def total_light_area(self) -> np.float32:
        return np.float32(self.lib.cupt_scene_total_light_area(self.h))

    def arrays(self):
        nv, nt, nm, nl, _ = self.counts()
        verts, tris = np.zeros(nv, dtype=VEC3), np.zeros(nt, dtype=TRIANGLE)
        mats, lights = np.zeros(nm, dtype=MATERIAL), np.zeros(nl, dtype=np.int32)
        _check(self.lib.cupt_scene_get_arrays(self.h, _ptr(verts), _ptr(tris), _ptr(mats), _ptr(lights)))
        return verts, tris, mats, lights

    def build_bvh2(self):
        """buildBVH() — BVH.h:443: returns (nodes[BVH2_NODE], depth)"""
        p, size, depth = C.c_void_p(), C.c_int(), C.c_int()
        _check(self.lib.cupt_bvh2_build(self.h, C.byref(p), C.byref(size), C.byref(depth)))
        try:
            buf = (C.c_char * (size.value * BVH2_NODE.itemsize)).from_address(p.value)
            nodes = np.frombuffer(buf, dtype=BVH2_NODE).copy()
        finally:
            self.lib.cupt_bvh2_free(p)
        return nodes, depth.value

    def build_bvh(self, sah_bins=0, max_leaf_tris=0, threads=0, tie_rank: Optional[np.ndarray] = None, greedy_collapse=False, reinsert_passes=0) -> "BvhHandle":
        return BvhHandle(self, sah_bins, max_leaf_tris, threads, tie_rank, greedy_collapse, reinsert_passes)


def bvh2_leaf_order(nodes: np.ndarray, ntris: int) -> np.ndarray:
    lib = load_library()
    nodes = np.ascontiguousarray(nodes, dtype=BVH2_NODE)
    rank = np.zeros(ntris, dtype=np.uint32)
    _check(lib.cupt_bvh2_leaf_order(_ptr(nodes), len(nodes), ntris, _ptr(rank)))
    return rank


NODE8 = np.dtype([("p", "<f4", (3,)), ("e", "u1", (3,)), ("imask", "u1"), ("child_base", "<u4"), ("reserved", "<u4"), ("tri_base", "<u4"),
                  ("pvalid", "<u4"), ("qlo", "u1", (3, 8)), ("qhi", "u1", (3, 8))])
PRIM48 = np.dtype([("a", "<f4", (3,)), ("b", "<f4", (3,)), ("c", "<f4", (3,)), ("id", "<u4"), ("rank", "<u4"), ("kind", "<u4")])
assert NODE8.itemsize == 80 and PRIM48.itemsize == 48


class BvhHandle:
    """cupt_bvh: host image of the 8-wide quantised BVH (replaces buildBVH, BVH.h:443-474)."""

    def __init__(self, scene: SceneHandle, sah_bins=0, max_leaf_tris=0, threads=0, tie_rank: Optional[np.ndarray] = None, greedy_collapse=False, reinsert_passes=0):
        self.h = None
        self.lib = load_library()
        self._rank = None if tie_rank is None else np.ascontiguousarray(tie_rank, dtype=np.uint32)
        opts = BvhOpts(sah_bins, max_leaf_tris, threads, None if self._rank is None else self._rank.ctypes.data, 0, int(reinsert_passes), int(greedy_collapse))
        h = C.c_void_p()
        _check(self.lib.cupt_bvh_build(scene.h, C.byref(opts), C.byref(h)))
        self.h = h

    def close(self):
        if self.h:
            self.lib.cupt_bvh_destroy(self.h)
            self.h = None

    __del__ = close

    def info(self) -> dict:
        i = BvhInfo()
        _check(self.lib.cupt_bvh_get_info(self.h, C.byref(i)))
        return {n: getattr(i, n) for n, _ in BvhInfo._fields_}

    def arrays(self):
        i = self.info()
        nodes = np.zeros(i["num_nodes"], dtype=NODE8)
        prims = np.zeros(i["num_tri_records"], dtype=PRIM48)
        _check(self.lib.cupt_bvh_get_arrays(self.h, _ptr(nodes), _ptr(prims)))
        return nodes, prims


class Context:
    """cupt_context: one GPU.  Mirrors the reference's main() flow (kernel.cu:633-761): set_scene / set_bvh /
    set_camera once, render(first_sample, n) any number of times, read_accum()."""

    def __init__(self, device: int = 0):
        self.h = None
        self.lib = load_library()
        h = C.c_void_p()
        _check(self.lib.cupt_create(device, C.byref(h)))
        self.h = h
        self.opts = RenderOpts()
        _check(self.lib.cupt_default_render_opts(C.byref(self.opts)))
        self._cam = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.cupt_destroy(self.h)
            self.h = None

    __del__ = close

    def set_stream(self, cuda_stream: int):
        _check(self.lib.cupt_set_stream(self.h, C.c_void_p(cuda_stream)))

    def set_scene(self, scene: SceneHandle):
        _check(self.lib.cupt_set_scene(self.h, scene.h))

    def set_bvh(self, bvh: BvhHandle):
        _check(self.lib.cupt_set_bvh(self.h, bvh.h))

    def build_bvh_device(self, max_leaf_tris=0, tie_rank: Optional[np.ndarray] = None, radius=0) -> dict:
        """cupt_build_bvh_device: wide BVH built on the GPU from the uploaded scene (replaces buildBVH + its upload)"""
        rank = None if tie_rank is None else np.ascontiguousarray(tie_rank, dtype=np.uint32)
        opts = BvhOpts(0, max_leaf_tris, 0, None if rank is None else rank.ctypes.data, radius, 0, 0)
        i = BvhInfo()
        _check(self.lib.cupt_build_bvh_device(self.h, C.byref(opts), C.byref(i)))
        d = {n: getattr(i, n) for n, _ in BvhInfo._fields_}
        ms = (C.c_float * 4)()
        _check(self.lib.cupt_get_build_times(self.h, ms))
        d.update(ms_total=ms[0], ms_sort=ms[1], ms_cluster=ms[2], ms_collapse=ms[3])
        return d

    def debug_sort_pairs(self, keys: np.ndarray, vals: np.ndarray, passes: int = 8):
        """test hook: the device builder's radix sort; returns sorted copies (stable, on the top 8 * passes of 63 bits)"""
        k = np.array(keys, dtype=np.uint64, copy=True); v = np.array(vals, dtype=np.uint32, copy=True)
        _check(self.lib.cupt_debug_sort_pairs(self.h, _ptr(k), _ptr(v), C.c_int64(len(k)), int(passes)))
        return k, v

    def release_workspace(self):
        _check(self.lib.cupt_release_workspace(self.h))

    def read_bvh(self):
        """(nodes, prims, depth, clustering rounds) of the wide BVH currently on the device"""
        counts = (C.c_int64 * 4)()
        _check(self.lib.cupt_read_bvh(self.h, None, None, counts))
        nodes = np.zeros(counts[0], dtype=NODE8)
        prims = np.zeros(counts[1], dtype=PRIM48)
        _check(self.lib.cupt_read_bvh(self.h, _ptr(nodes), _ptr(prims), counts))
        return nodes, prims, int(counts[2]), int(counts[3])

    def set_bvh2(self, nodes: np.ndarray, depth: int):
        nodes = np.ascontiguousarray(nodes, dtype=BVH2_NODE)
        _check(self.lib.cupt_set_bvh2(self.h, _ptr(nodes), len(nodes), depth))

    def set_camera(self, cam: np.ndarray):
        self._cam = np.ascontiguousarray(cam, dtype=CAMERA)
        _check(self.lib.cupt_set_camera(self.h, _ptr(self._cam)))

    def set_options(self, **kw):
        for k, v in kw.items():
            if not hasattr(self.opts, k):
                raise AttributeError(k)
            setattr(self.opts, k, v)
        _check(self.lib.cupt_set_render_opts(self.h, C.byref(self.opts)))

    @property
    def shape(self):
        return int(self._cam["pxlHeight"][0]), int(self._cam["pxlWidth"][0])

    def render(self, first_sample: int, num_samples: int):
        _check(self.lib.cupt_render(self.h, first_sample, num_samples))

    def sync(self):
        _check(self.lib.cupt_sync(self.h))

    def set_tile_window(self, first_tile: int = 0, num_tiles: int = 0):
        """render only tiles [first_tile, first_tile + num_tiles) of 8x4 pixels, row-major (0 tiles = whole image): the
        image-tile partition of the multi-GPU path (sharding.tile_range)"""
        _check(self.lib.cupt_set_tile_window(self.h, int(first_tile), int(num_tiles)))

    def tile_count(self) -> int:
        n = C.c_int(0)
        _check(self.lib.cupt_tile_count(self.h, C.byref(n)))
        return n.value

    def reset_accum(self):
        _check(self.lib.cupt_reset_accum(self.h))

    def read_accum(self, normalise: bool = True, out: Optional[np.ndarray] = None) -> np.ndarray:
        h, w = self.shape
        if out is None:
            out = np.empty((h * w, 3), dtype=np.float32)
        _check(self.lib.cupt_read_accum(self.h, _ptr(out), int(normalise)))
        return out

    def read_image_rgb8(self, mirror_x: bool = False) -> np.ndarray:
        """tonemapped 8-bit image [H, W, 3] computed on the device (kernel.cu:763-778 without the file)"""
        h, w = self.shape
        out = np.empty((h, w, 3), dtype=np.uint8)
        _check(self.lib.cupt_read_image_rgb8(self.h, _ptr(out), int(mirror_x)))
        return out

    def read_heatmap(self):
        """(nodes visited, primitives tested) per pixel by the primary rays, two [H, W] float arrays"""
        h, w = self.shape
        a = np.empty((h, w), dtype=np.float32); b = np.empty((h, w), dtype=np.float32)
        _check(self.lib.cupt_read_heatmap(self.h, _ptr(a), _ptr(b)))
        return a, b

    def copy_accum_device(self, dst_ptr: int):
        _check(self.lib.cupt_copy_accum_device(self.h, C.c_void_p(dst_ptr)))

    def add_accum_device(self, src_ptr: int, samples: int):
        _check(self.lib.cupt_add_accum_device(self.h, C.c_void_p(src_ptr), samples))

    def primary_hits(self):
        h, w = self.shape
        ids, ts = np.empty(h * w, dtype=np.int32), np.empty(h * w, dtype=np.float32)
        _check(self.lib.cupt_read_primary_hits(self.h, _ptr(ids), _ptr(ts)))
        return ids, ts

    def read_retrace_counts(self) -> np.ndarray:
        """strict mode with record_retraces=1: per pixel, the rays since reset_accum whose wide hit the reference would not have reached"""
        h, w = self.shape
        out = np.empty(h * w, dtype=np.uint32)
        _check(self.lib.cupt_read_retrace_counts(self.h, _ptr(out)))
        return out

    def camera_rays(self):
        h, w = self.shape
        rays = np.empty((h * w, 6), dtype=np.float32)
        _check(self.lib.cupt_camera_rays(self.h, _ptr(rays)))
        return rays

    def trace_rays(self, rays: np.ndarray):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        ids, ts = np.empty(len(rays), dtype=np.int32), np.empty(len(rays), dtype=np.float32)
        _check(self.lib.cupt_trace_rays(self.h, _ptr(rays), C.c_int64(len(rays)), _ptr(ids), _ptr(ts)))
        return ids, ts

    def trace_rays_noalloc(self, rays: np.ndarray):
        """cupt_trace_rays_noalloc on host arrays: same hits as trace_rays, no allocation inside the library (<= NOALLOC_MAX_RAYS rays)"""
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 6)
        ids, ts = np.empty(len(rays), dtype=np.int32), np.empty(len(rays), dtype=np.float32)
        _check(self.lib.cupt_trace_rays_noalloc(self.h, _ptr(rays), C.c_int64(len(rays)), _ptr(ids), _ptr(ts), 1))
        return ids, ts

    def trace_rays_noalloc_ptr(self, rays_ptr: int, n: int, ids_ptr: int, ts_ptr: int, synchronise: bool = True):
        """the same on raw (host or device) addresses"""
        _check(self.lib.cupt_trace_rays_noalloc(self.h, C.c_void_p(rays_ptr), C.c_int64(n), C.c_void_p(ids_ptr), C.c_void_p(ts_ptr), int(synchronise)))

    def stats(self) -> dict:
        s = Stats()
        _check(self.lib.cupt_get_stats(self.h, C.byref(s)))
        return s.as_dict()

    def reset_stats(self):
        _check(self.lib.cupt_reset_stats(self.h))


class _BorrowedContext(Context):
    """a group member's context: owned (and destroyed) by the group"""

    def __init__(self, handle, cam):
        self.h, self.lib, self.opts, self._cam = handle, load_library(), RenderOpts(), cam
        _check(self.lib.cupt_default_render_opts(C.byref(self.opts)))

    def close(self):
        self.h = None

    __del__ = close


def group_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    _check(load_library().cupt_group_unique_id(buf))
    return buf.raw


class Group:
    """cupt_group: several GPUs behind one handle.  Group(devices=[0, 1, ...]) drives every member from this process (one
    host thread per GPU); Group(device=d, rank=r, nranks=n, unique_id=id) is this process's member of an n-process group."""

    def __init__(self, devices=None, device=None, rank=None, nranks=None, unique_id=None):
        self.h = None
        self.lib = load_library()
        self.lib.cupt_group_context.restype = C.c_void_p
        self.lib.cupt_group_context.argtypes = [C.c_void_p, C.c_int]
        self.lib.cupt_group_destroy.restype = None
        self.lib.cupt_group_destroy.argtypes = [C.c_void_p]
        h = C.c_void_p()
        if devices is not None:
            arr = (C.c_int * len(devices))(*devices)
            _check(self.lib.cupt_group_create_local(arr, len(devices), C.byref(h)))
        else:
            _check(self.lib.cupt_group_create_rank(int(device), int(rank), int(nranks), C.c_char_p(unique_id), C.byref(h)))
        self.h = h
        self.opts = RenderOpts()
        _check(self.lib.cupt_default_render_opts(C.byref(self.opts)))
        self._cam = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.cupt_group_destroy(self.h)
            self.h = None

    __del__ = close

    def info(self):
        out = (C.c_int * 3)()
        _check(self.lib.cupt_group_info(self.h, out))
        return {"nranks": out[0], "nlocal": out[1], "first_rank": out[2]}

    def context(self, local_index: int) -> Context:
        p = self.lib.cupt_group_context(self.h, local_index)
        if not p:
            _check(-1)
        return _BorrowedContext(C.c_void_p(p), self._cam)

    def set_scene(self, scene: Optional[SceneHandle]):
        _check(self.lib.cupt_group_set_scene(self.h, None if scene is None else scene.h))

    def set_bvh(self, bvh: Optional[BvhHandle]):
        _check(self.lib.cupt_group_set_bvh(self.h, None if bvh is None else bvh.h))

    def build_bvh_device(self, max_leaf_tris=0, tie_rank: Optional[np.ndarray] = None, radius=0) -> dict:
        rank = None if tie_rank is None else np.ascontiguousarray(tie_rank, dtype=np.uint32)
        opts = BvhOpts(0, max_leaf_tris, 0, None if rank is None else rank.ctypes.data, radius, 0, 0)
        i = BvhInfo()
        _check(self.lib.cupt_group_build_bvh_device(self.h, C.byref(opts), C.byref(i)))
        return {n: getattr(i, n) for n, _ in BvhInfo._fields_}

    def set_bvh2(self, nodes: Optional[np.ndarray], depth: int = 0):
        if nodes is None:
            _check(self.lib.cupt_group_set_bvh2(self.h, None, 0, 0))
            return
        nodes = np.ascontiguousarray(nodes, dtype=BVH2_NODE)
        _check(self.lib.cupt_group_set_bvh2(self.h, _ptr(nodes), len(nodes), depth))

    def set_camera(self, cam: np.ndarray):
        self._cam = np.ascontiguousarray(cam, dtype=CAMERA)
        _check(self.lib.cupt_group_set_camera(self.h, _ptr(self._cam)))

    def set_options(self, **kw):
        for k, v in kw.items():
            if not hasattr(self.opts, k):
                raise AttributeError(k)
            setattr(self.opts, k, v)
        _check(self.lib.cupt_group_set_render_opts(self.h, C.byref(self.opts)))

    def reset_accum(self):
        _check(self.lib.cupt_group_reset_accum(self.h))

    def render(self, first_sample: int, num_samples: int):
        _check(self.lib.cupt_group_render(self.h, first_sample, num_samples))

    def sync(self):
        _check(self.lib.cupt_group_sync(self.h))

    def read_accum(self, normalise: bool = True, out: Optional[np.ndarray] = None):
        """the image from rank 0; None on a process that does not hold rank 0"""
        h, w = int(self._cam["pxlHeight"][0]), int(self._cam["pxlWidth"][0])
        if out is None:
            out = np.empty((h * w, 3), dtype=np.float32)
        rc = self.lib.cupt_group_read_accum(self.h, _ptr(out), int(normalise))
        if rc == 1:
            return None
        _check(rc)
        return out

    def read_image_rgb8(self, mirror_x: bool = False):
        h, w = int(self._cam["pxlHeight"][0]), int(self._cam["pxlWidth"][0])
        out = np.empty((h, w, 3), dtype=np.uint8)
        rc = self.lib.cupt_group_read_image_rgb8(self.h, _ptr(out), int(mirror_x))
        if rc == 1:
            return None
        _check(rc)
        return out

    def stats(self) -> dict:
        s = Stats()
        _check(self.lib.cupt_group_get_stats(self.h, C.byref(s)))
        return s.as_dict()


def write_ppm(path: str, rgb_mean: np.ndarray, width: int, height: int):
    """kernel.cu:763-778"""
    rgb_mean = np.ascontiguousarray(rgb_mean, dtype=np.float32)
    _check(load_library().cupt_write_ppm(path.encode(), _ptr(rgb_mean), width, height))


def write_pfm(path: str, rgb: np.ndarray, width: int, height: int):
    rgb = np.ascontiguousarray(rgb, dtype=np.float32)
    assert rgb.size == width * height * 3
    _check(load_library().cupt_write_pfm(path.encode(), _ptr(rgb), width, height))


def write_ppm8(path: str, rgb8: np.ndarray):
    rgb8 = np.ascontiguousarray(rgb8, dtype=np.uint8)
    _check(load_library().cupt_write_ppm8(path.encode(), _ptr(rgb8), rgb8.shape[1], rgb8.shape[0]))


def tonemap(rgb_mean: np.ndarray) -> np.ndarray:
    rgb_mean = np.ascontiguousarray(rgb_mean, dtype=np.float32).reshape(-1, 3)
    out = np.empty((len(rgb_mean), 3), dtype=np.int32)
    _check(load_library().cupt_tonemap(_ptr(rgb_mean), C.c_int64(len(rgb_mean)), _ptr(out)))
    return out
