"""TEST / BASELINE INFRASTRUCTURE: ctypes binding of oracle/_ref/libcupt_legacy.so — the reference's kernel.cu rebuilt
for sm_100a with its compile patches only (oracle/build_ref.sh).  Runs on a GPU box; used by bench.py's
`legacy_gpu` leg and tools/, never by the product."""
import ctypes as C
import os

import numpy as np

PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "libcupt_legacy.so")
_lib = None


def available():
    return os.path.exists(PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(PATH)
    return _lib


def image_dim():
    return int(lib().legacy_image_dim())


def render(verts, tris, mats, lights, area, cam, num_samples, want_image=False):
    """drawPixel loop of kernel.cu:702-739 on the GPU.  Returns dict(ms, rays, build_s, depth, image)."""
    L = lib()
    dim = image_dim()
    img = np.zeros((dim * dim, 3), np.float64) if want_image else None
    ms, rays, bs, depth = C.c_float(), C.c_ulonglong(), C.c_double(), C.c_int()
    pos = (C.c_float * 3)(*[float(x) for x in cam["pos"][0]])
    P = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
    v = np.ascontiguousarray(verts); t = np.ascontiguousarray(tris); m = np.ascontiguousarray(mats)
    l = np.ascontiguousarray(lights, dtype=np.int32)
    devnull = os.open(os.devnull, os.O_WRONLY); saved = os.dup(1); os.dup2(devnull, 1)   # buildBVH() printf()s
    try:
        rc = L.legacy_gpu_render(P(v), len(v), P(t), len(t), P(m), len(m), P(l), len(l), C.c_float(float(area)), pos,
                                 C.c_float(float(cam["distFromFilm"][0])), C.c_float(float(cam["focalLength"][0])),
                                 C.c_float(float(cam["radius"][0])), int(num_samples), P(img), C.byref(ms), C.byref(rays),
                                 C.byref(bs), C.byref(depth))
    finally:
        os.dup2(saved, 1); os.close(devnull); os.close(saved)
    if rc != 0:
        raise RuntimeError(f"legacy_gpu_render failed: {rc} (bvh depth {depth.value})")
    return {"ms": ms.value, "rays": rays.value, "build_s": bs.value, "depth": depth.value, "image": img}
