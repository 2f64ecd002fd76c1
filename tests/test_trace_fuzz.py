"""Differential fuzz of trace() (kernel.cu:112-161, with rayAABBIntersect BVH.h:51-83 and triIntersect modelLoader.h:49-82):
the REAL reference compiled into oracle/_ref, the oracle's C restatement and the product's reference-mode traversal header
(traverse_bvh2_ref in csrc/traverse.cuh, compiled for the host by tests/emul) must return the same triangle and the same bits of t on
every ray of hypothesis-generated scenes — including the rays on which the slab test divides 0 by 0 (axis-parallel rays through box
planes), rays from and through vertices, unnormalised directions, degenerate triangles.  This is the oracle's pin on inputs the golden
cases do not contain.  Skips where /root/reference was not available at build time (fixture `ref`)."""
import ctypes as C

import os

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from cudapathtracer_b200 import api
from test_bvh8 import P, emul  # noqa: F401  (emul: fixture)
from test_bvh8_property import make_rays, tiny_scenes


@pytest.mark.timeout(600)
@settings(max_examples=int(os.environ.get("CUPT_FUZZ_EXAMPLES", "200")), deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow], derandomize=True)
@given(tiny_scenes())
def test_reference_trace_oracle_and_product_header_agree(lib, ref, oracle, emul, scene):
    v, idx, seed = scene
    if len(idx) < 2:
        return
    verts = np.ascontiguousarray(v).view(api.VEC3).reshape(-1)
    tris = np.zeros(len(idx), dtype=api.TRIANGLE)
    tris["v0"], tris["v1"], tris["v2"] = idx[:, 0], idx[:, 1], idx[:, 2]
    mats = np.zeros(1, dtype=api.MATERIAL)
    s = api.SceneHandle()
    s.set_arrays(verts, tris, mats, np.zeros(0, np.int32), 0.0)
    sv, stt, sm, sl = s.arrays()
    try:
        n2, d2 = s.build_bvh2()
    except api.CuptError:
        s.close()
        return                                                                  # the reference cannot build this one (tests/test_bvh2_fuzz.py)
    rays = make_rays(v, idx, seed)
    ref.reset()
    ref.set_arrays(sv, stt, sm, sl, 0.0)
    rn, rd = ref.build_bvh()
    assert rn.tobytes() == n2.tobytes()
    ri, rt = ref.trace_rays(rays)                                               # the real trace()
    osc = oracle.OracleScene(sv, stt, sm, sl, 0.0, n2)
    oi, ot = osc.trace_rays(rays)                                               # the oracle's restatement
    assert np.array_equal(oi, ri) and ot.tobytes() == rt.tobytes(), np.flatnonzero((oi != ri) | (ot.view(np.uint32) != rt.view(np.uint32)))[:5]
    ids, ts = np.zeros(len(rays), np.int32), np.zeros(len(rays), np.float32)
    vf = np.ascontiguousarray(sv.view(np.float32))
    emul.emul_trace_bvh2(P(n2), P(vf), P(stt), P(rays), C.c_longlong(len(rays)), P(ids), P(ts))   # the product's header
    assert np.array_equal(ids, ri) and ts.tobytes() == rt.tobytes(), np.flatnonzero((ids != ri) | (ts.view(np.uint32) != rt.view(np.uint32)))[:5]
    s.close()
