"""The property of tests/test_bvh8_property.py on the GPU, through the C-ABI: for any (small, nasty) scene and any ray without
degenerate-triangle noise, cupt_trace_rays returns the oracle's brute-force argmin over all triangles bit for bit — through the
host-built tree and through the tree built on the device, wide and strict modes being out of the picture here (no reference tree:
its builder rejects deep trees of coincident triangles, tests/test_bvh2.py)."""
import os

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from cudapathtracer_b200 import api
from test_bvh8_property import make_rays, tiny_scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx(lib):
    c = api.Context(0)
    yield c
    c.close()


@settings(max_examples=int(os.environ.get("CUPT_FUZZ_EXAMPLES", "600")), deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow], derandomize=True)
@given(tiny_scenes(), st.sampled_from([1, 3]))
def test_gpu_traversal_equals_brute_force_on_any_scene(ctx, oracle, scene, max_leaf):
    v, idx, seed = scene
    verts = np.ascontiguousarray(v).view(api.VEC3).reshape(-1)
    tris = np.zeros(len(idx), dtype=api.TRIANGLE)
    tris["v0"], tris["v1"], tris["v2"] = idx[:, 0], idx[:, 1], idx[:, 2]
    tris["norm"] = (0, 1, 0)
    mats = np.zeros(1, dtype=api.MATERIAL); mats["albedo"] = (0.5, 0.5, 0.5)
    s = api.SceneHandle()
    s.set_arrays(verts, tris, mats, np.zeros(0, np.int32), 0.0)
    sv, stt, sm, sl = s.arrays()
    rays = make_rays(v, idx, seed)
    osc = oracle.OracleScene(sv, stt, sm, sl, 0.0, np.zeros(1, api.BVH2_NODE))
    want_i, want_t = osc.brute_force(rays, None)
    noisy = osc.hits_outside_box(rays)
    ctx.set_scene(s)
    ctx.set_options(traversal=api.TRAVERSAL_WIDE, tri_test=api.TRI_REFEXACT)
    for builder in ("host", "device"):
        if builder == "host":
            b = s.build_bvh(max_leaf_tris=max_leaf)
            ctx.set_bvh(b)
        else:
            b = None
            ctx.build_bvh_device(max_leaf_tris=max_leaf)
        ids, ts = ctx.trace_rays(rays)
        same = (ids == want_i) & (ts.view(np.uint32) == want_t.view(np.uint32))
        bad = np.flatnonzero(~same & ~noisy)
        assert len(bad) == 0, (builder, bad[:5], ids[bad[:5]], want_i[bad[:5]], ts[bad[:5]], want_t[bad[:5]], rays[bad[:5]])
        assert np.all(ts[noisy] >= want_t[noisy])
        if b is not None:
            b.close()
    s.close()


@settings(max_examples=int(os.environ.get("CUPT_FUZZ_EXAMPLES", "600")), deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow], derandomize=True)
@given(tiny_scenes())
def test_gpu_reference_modes_equal_trace_on_any_scene(ctx, oracle, scene):
    """REF_BVH2 mode == the oracle's trace() (pinned to the real reference on the same generator by tests/test_trace_fuzz.py) on EVERY
    ray, 0/0 in the slab test and degenerate triangles included; STRICT mode == trace() on every ray without degenerate-triangle noise
    (DESIGN.md section 2), and on those never nearer than it."""
    v, idx, seed = scene
    if len(idx) < 2:
        return
    verts = np.ascontiguousarray(v).view(api.VEC3).reshape(-1)
    tris = np.zeros(len(idx), dtype=api.TRIANGLE)
    tris["v0"], tris["v1"], tris["v2"] = idx[:, 0], idx[:, 1], idx[:, 2]
    tris["norm"] = (0, 1, 0)
    mats = np.zeros(1, dtype=api.MATERIAL)
    s = api.SceneHandle()
    s.set_arrays(verts, tris, mats, np.zeros(0, np.int32), 0.0)
    sv, stt, sm, sl = s.arrays()
    try:
        n2, d2 = s.build_bvh2()
    except api.CuptError:
        s.close()
        return
    rank = api.bvh2_leaf_order(n2, len(stt))
    rays = make_rays(v, idx, seed)
    osc = oracle.OracleScene(sv, stt, sm, sl, 0.0, n2)
    want_i, want_t = osc.trace_rays(rays)
    noisy = osc.hits_outside_box(rays)
    b = s.build_bvh(tie_rank=rank)
    ctx.set_scene(s); ctx.set_bvh(b); ctx.set_bvh2(n2, d2)
    ctx.set_options(traversal=api.TRAVERSAL_REF_BVH2, tri_test=api.TRI_REFEXACT)
    ids, ts = ctx.trace_rays(rays)
    bad = np.flatnonzero((ids != want_i) | (ts.view(np.uint32) != want_t.view(np.uint32)))
    assert len(bad) == 0, ("ref_bvh2", bad[:5], ids[bad[:5]], want_i[bad[:5]], ts[bad[:5]], want_t[bad[:5]], rays[bad[:5]])
    ctx.set_options(traversal=api.TRAVERSAL_WIDE_STRICT, tri_test=api.TRI_REFEXACT)
    ids, ts = ctx.trace_rays(rays)
    same = (ids == want_i) & (ts.view(np.uint32) == want_t.view(np.uint32))
    bad = np.flatnonzero(~same & ~noisy)
    assert len(bad) == 0, ("strict", bad[:5], ids[bad[:5]], want_i[bad[:5]], ts[bad[:5]], want_t[bad[:5]], rays[bad[:5]])
    b.close(); s.close()
