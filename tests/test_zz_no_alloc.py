"""cupt_trace_rays_noalloc — the allocation-free variant of cupt_trace_rays (no counterpart in the reference; added
for callers that cannot call cudaMalloc).  Its scratch is a fixed file-scope device array,
shared by every call and never cleared between calls except for what must start at zero, so the test makes TWO calls in a
row with different ray sets of different sizes and compares each with the existing entry point: hits are bit-exact
(ids equal, t bits equal), the tolerance every hit test in this suite uses.

Collected last on purpose; it builds its own context and scene, so it passes alone in a
fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from cudapathtracer_b200 import api, scenes

pytestmark = pytest.mark.gpu


def random_rays(lo, hi, n, seed):
    rng = np.random.default_rng(seed)
    o = rng.uniform(lo, hi, size=(n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d], axis=1).astype(np.float32)


def same(a, b):
    return np.array_equal(a[0], b[0]) and a[1].tobytes() == b[1].tobytes()


def test_no_alloc_trace_equals_trace_rays_twice_in_a_row(lib):
    sc = scenes.atrium(20000)
    s = api.SceneHandle(); s.add_scene(sc)
    n2, d2 = s.build_bvh2()
    bvh = s.build_bvh(tie_rank=api.bvh2_leaf_order(n2, s.counts()[1]))
    v = s.arrays()[0].view(np.float32).reshape(-1, 3)
    ctx = api.Context(0)
    try:
        ctx.set_scene(s); ctx.set_bvh(bvh); ctx.set_bvh2(n2, d2)
        ctx.set_options(traversal=api.TRAVERSAL_WIDE)
        first = random_rays(v.min(0), v.max(0), 50000, 1)
        second = random_rays(v.min(0), v.max(0), 777, 2)          # smaller: the array still holds the first call's records beyond 777
        want1, want2 = ctx.trace_rays(first), ctx.trace_rays(second)
        assert (want1[0] >= 0).mean() > 0.9
        got1 = ctx.trace_rays_noalloc(first)
        got2 = ctx.trace_rays_noalloc(second)
        assert same(got1, want1)
        assert same(got2, want2)
        assert same(ctx.trace_rays_noalloc(first), want1)          # and the large batch again after the small one
        # the reference's own algorithm through the same entry point
        ctx.set_options(traversal=api.TRAVERSAL_REF_BVH2)
        assert same(ctx.trace_rays_noalloc(second), ctx.trace_rays(second))
        ctx.set_options(traversal=api.TRAVERSAL_WIDE)
        # limits and refusals: a clear error, nothing traced
        ids, ts = ctx.trace_rays_noalloc(np.zeros((0, 6), np.float32))
        assert len(ids) == 0
        full = random_rays(v.min(0), v.max(0), api.NOALLOC_MAX_RAYS, 3)
        assert same(ctx.trace_rays_noalloc(full), ctx.trace_rays(full))
        with pytest.raises(api.CuptError, match="CUPT_NOALLOC_MAX_RAYS"):
            ctx.trace_rays_noalloc(np.zeros((api.NOALLOC_MAX_RAYS + 1, 6), np.float32))
        ctx.set_options(traversal=api.TRAVERSAL_WIDE_STRICT)
        with pytest.raises(api.CuptError, match="strict"):
            ctx.trace_rays_noalloc(second)
        ctx.set_options(traversal=api.TRAVERSAL_WIDE)
        # device pointers in and out (the caller's memory: allocated here with the CUDA runtime, not by the library)
        rt = C.CDLL("libcudart.so.12")
        n = len(second)
        d_r, d_i, d_t = C.c_void_p(), C.c_void_p(), C.c_void_p()
        assert rt.cudaMalloc(C.byref(d_r), C.c_size_t(n * 24)) == 0 and rt.cudaMalloc(C.byref(d_i), C.c_size_t(n * 4)) == 0 and rt.cudaMalloc(C.byref(d_t), C.c_size_t(n * 4)) == 0
        try:
            assert rt.cudaMemcpy(d_r, second.ctypes.data_as(C.c_void_p), C.c_size_t(n * 24), 1) == 0
            ctx.trace_rays_noalloc_ptr(d_r.value, n, d_i.value, d_t.value, True)
            ids, ts = np.empty(n, np.int32), np.empty(n, np.float32)
            assert rt.cudaMemcpy(ids.ctypes.data_as(C.c_void_p), d_i, C.c_size_t(n * 4), 2) == 0
            assert rt.cudaMemcpy(ts.ctypes.data_as(C.c_void_p), d_t, C.c_size_t(n * 4), 2) == 0
            assert same((ids, ts), want2)
        finally:
            rt.cudaFree(d_r); rt.cudaFree(d_i); rt.cudaFree(d_t)
    finally:
        ctx.close()
