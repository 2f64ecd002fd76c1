"""Differential fuzz of the reference-layout BVH2 builder: cupt_bvh2_build (csrc/bvh2_compat.cpp) against the REAL buildBVH()
(BVH.h:443-474, compiled into oracle/_ref) on hypothesis-generated scenes — lattice vertices (coincident centroids, identical boxes:
the positional-halving path BVH.h:263-288), duplicated, collinear and point triangles, one huge triangle among tiny ones, coordinates
at the edge of the +-1e4 world (BVH.h:21-25).  Node arrays and depth must be equal byte for byte; the leaf order derived from them is
the tie-break rank the wide BVH carries.  Skips where /root/reference was not available at build time (fixture `ref`)."""
import os

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from cudapathtracer_b200 import api
from test_bvh8_property import tiny_scenes


@pytest.mark.timeout(600)
@settings(max_examples=int(os.environ.get("CUPT_FUZZ_EXAMPLES", "200")), deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow], derandomize=True)
@given(tiny_scenes(), st.integers(1, 6))
def test_generated_scenes_build_the_reference_tree(lib, ref, scene, copies):
    v, idx, seed = scene
    if len(idx) < 2:
        return
    rng = np.random.default_rng(seed)
    # several shifted copies: more triangles (up to ~290), repeated structure
    vs, ids = [], []
    for c in range(copies):
        shift = (rng.integers(-2, 3, size=3) * np.float32(np.abs(v).max() + 1)).astype(np.float32) if c else np.zeros(3, np.float32)
        vs.append(v + shift); ids.append(idx + c * len(v))
    v, idx = np.concatenate(vs), np.concatenate(ids)
    verts = np.ascontiguousarray(v).view(api.VEC3).reshape(-1)
    tris = np.zeros(len(idx), dtype=api.TRIANGLE)
    tris["v0"], tris["v1"], tris["v2"] = idx[:, 0], idx[:, 1], idx[:, 2]
    mats = np.zeros(1, dtype=api.MATERIAL)
    s = api.SceneHandle()
    s.set_arrays(verts, tris, mats, np.zeros(0, np.int32), 0.0)
    sv, stt, sm, sl = s.arrays()
    # the product first: on input the reference's recursion never leaves (an empty split side and nothing to halve, BVH.h:263-276 —
    # the real buildBVH() overflows its stack and takes the process with it) it returns an error code, and the reference is not called
    try:
        n, d = s.build_bvh2()
    except api.CuptError as e:
        assert "does not terminate" in str(e) or "deeper than" in str(e), str(e)
        s.close()
        return
    ref.reset()
    ref.set_arrays(sv, stt, sm, sl, 0.0)
    rn, rd = ref.build_bvh()
    assert d == rd and len(n) == len(rn), (d, rd, len(n), len(rn))
    assert n.tobytes() == rn.tobytes()
    order = api.bvh2_leaf_order(n, len(stt))
    assert np.array_equal(np.sort(order), np.arange(len(stt), dtype=order.dtype))     # a permutation: every triangle is a leaf once
    s.close()
