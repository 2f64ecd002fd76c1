"""The property of tests/test_bvh8_property.py on the GPU, through the C-ABI: for any (small, nasty) scene and any ray without
degenerate-triangle noise, cupt_trace_rays returns the oracle's brute-force argmin over all triangles bit for bit — through the
host-built tree and through the tree built on the device, wide and strict modes being out of the picture here (no reference tree:
its builder rejects deep trees of coincident triangles, tests/test_bvh2.py)."""
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


@settings(max_examples=150, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow], derandomize=True)
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
