"""Property test of the wide BVH + traversal logic on the CPU (host builder + tests/emul, the header k_extend compiles):

    for ANY scene and ANY ray, the wide traversal returns the argmin over all triangles of triIntersect's t with 0 < t < 1e5
    (modelLoader.h:49-82, kernel.cu:117,127), ties by rank — bit for bit what the oracle's brute force returns.

hypothesis draws small scenes built to provoke the edge cases the fixed parity cases only touch by accident: vertices on a coarse
lattice (coplanar and coincident triangles, exact ties in t), duplicated triangles, zero-area triangles, a huge triangle among tiny
ones, coordinates at the edge of the reference's +-1e4 world (BVH.h:21-25); rays from lattice points through lattice points (hits on
edges and vertices), axis-parallel rays (1/0 in the slab test), rays starting on a triangle's plane."""
import ctypes as C

import os

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from cudapathtracer_b200 import api
from test_bvh8 import P, check_wide_bvh, emul, emul_trace  # noqa: F401  (emul: fixture)


class Tiny:
    """what check_wide_bvh reads of a parity case"""
    def __init__(self, verts, tris):
        self.verts, self.tris, self.spheres = verts, tris, None
        self.rank = np.arange(len(tris), dtype=np.uint32)                       # no tie_rank given: ties by triangle index


@st.composite
def tiny_scenes(draw):
    nt = draw(st.integers(1, 48))
    scale = draw(st.sampled_from([1.0, 1.0, 0.125, 37.5, 2500.0]))          # 2500 * 4 = the +-1e4 world
    seed = draw(st.integers(0, 2 ** 32 - 1))
    rng = np.random.default_rng(seed)
    lattice = draw(st.booleans())
    if lattice:
        v = rng.integers(-4, 5, size=(3 * nt, 3)).astype(np.float32)
    else:
        v = rng.uniform(-4, 4, size=(3 * nt, 3)).astype(np.float32)
    v *= np.float32(scale)
    idx = np.arange(3 * nt, dtype=np.int32).reshape(nt, 3)
    kind = draw(st.lists(st.sampled_from(["plain", "plain", "plain", "dup", "flat", "point", "huge", "sliver"]), min_size=nt, max_size=nt))
    for i, k in enumerate(kind):
        a, b, c = idx[i]
        if k == "dup" and i > 0:
            idx[i] = idx[rng.integers(0, i)]                                     # the same three vertices again: every hit is a tie
        elif k == "flat":
            v[c] = v[a] + (v[b] - v[a]) * np.float32(0.5)                        # zero area (collinear)
        elif k == "point":
            v[b] = v[a]; v[c] = v[a]
        elif k == "huge":
            v[a] = np.float32(scale) * np.array([-4, -4, rng.integers(-4, 5)], np.float32)
            v[b] = np.float32(scale) * np.array([4, -4, v[a][2] / scale], np.float32)
            v[c] = np.float32(scale) * np.array([0, 4, v[a][2] / scale], np.float32)
        elif k == "sliver":
            v[c] = v[a] + (v[b] - v[a]) * np.float32(0.5) + np.float32(scale * 1e-4) * rng.normal(size=3).astype(np.float32)
    return v, idx, seed


def make_rays(v, idx, seed, n=160):
    rng = np.random.default_rng(seed ^ 0x5bd1e995)
    lo, hi = v.min(0), v.max(0)
    ext = np.maximum(hi - lo, 1e-3)
    rays = []
    # (1) random
    o = rng.uniform(lo - 0.5 * ext, hi + 0.5 * ext, size=(n, 3)); d = rng.normal(size=(n, 3))
    rays.append(np.concatenate([o, d / np.linalg.norm(d, axis=1, keepdims=True)], 1))
    # (2) from a vertex-like point through a vertex / an edge midpoint / a centroid (not normalised: the reference never requires it)
    tri = v[idx]                                                              # (nt, 3, 3)
    pick = rng.integers(0, len(idx), n)
    w = rng.choice([0.0, 0.5, 1.0, 1 / 3], size=(n, 3)); w[:, 2] = 1 - w[:, 0] - w[:, 1]
    target = np.einsum("nk,nkc->nc", w, tri[pick].astype(np.float64))
    o = v[rng.integers(0, len(v), n)].astype(np.float64) + rng.choice([0.0, 1.0, -2.0], size=(n, 3)) * ext
    d = target - o
    keep = np.abs(d).sum(1) > 0
    rays.append(np.concatenate([o[keep], d[keep]], 1))
    # (3) axis-parallel, through a vertex
    o = v[rng.integers(0, len(v), n // 4)].astype(np.float64)
    d = np.zeros((n // 4, 3)); ax = rng.integers(0, 3, n // 4); d[np.arange(n // 4), ax] = rng.choice([-1.0, 1.0], n // 4)
    o = o - d * ext[ax][:, None] * 2
    rays.append(np.concatenate([o, d], 1))
    # (4) starting ON a triangle's plane (t = 0 is not a hit, kernel.cu:127)
    o = target[: n // 4]; d = rng.normal(size=(n // 4, 3))
    rays.append(np.concatenate([o, d], 1))
    return np.concatenate(rays).astype(np.float32)


@settings(max_examples=int(os.environ.get("CUPT_FUZZ_EXAMPLES", "300")), deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow], derandomize=True)
@given(tiny_scenes(), st.sampled_from([1, 3]), st.sampled_from([-1, 0]))
def test_wide_traversal_equals_brute_force_on_any_scene(lib, oracle, emul, scene, max_leaf, reinsert):
    v, idx, seed = scene
    verts = np.ascontiguousarray(v).view(api.VEC3).reshape(-1)
    tris = np.zeros(len(idx), dtype=api.TRIANGLE)
    tris["v0"], tris["v1"], tris["v2"] = idx[:, 0], idx[:, 1], idx[:, 2]
    tris["norm"] = (0, 1, 0)
    mats = np.zeros(1, dtype=api.MATERIAL); mats["albedo"] = (0.5, 0.5, 0.5)
    s = api.SceneHandle()
    s.set_arrays(verts, tris, mats, np.zeros(0, np.int32), 0.0)
    b = s.build_bvh(max_leaf_tris=max_leaf, reinsert_passes=reinsert)           # ties by triangle index (no reference tree: its
    info = b.info()                                                             # builder rejects deep trees, tests/test_bvh2.py)
    nodes, prims = b.arrays()
    sv, stt, sm, sl = s.arrays()
    check_wide_bvh(Tiny(sv, stt), nodes, prims, info, max_leaf)
    rays = make_rays(v, idx, seed)
    osc = oracle.OracleScene(sv, stt, sm, sl, 0.0, np.zeros(1, api.BVH2_NODE))
    want_i, want_t = osc.brute_force(rays, None)
    # Rays on which some triangle "hits" outside its own bounding box are left out of the equality: that is rounding noise of a
    # numerically degenerate triangle (oracle.c: orc_hits_outside_box), which any tree — the reference's too — finds or not depending
    # on which boxes the ray happens to pass.  There the traversal may only return a hit that exists, at or behind the brute force's.
    noisy = osc.hits_outside_box(rays)
    for fast in (False, True):                                                   # the definition, and the node step k_extend runs
        ids, ts, _ = emul_trace(emul, nodes, prims, info["depth"], rays, fast=fast)
        same = (ids == want_i) & (ts.view(np.uint32) == want_t.view(np.uint32))
        bad = np.flatnonzero(~same & ~noisy)
        assert len(bad) == 0, (fast, bad[:5], ids[bad[:5]], want_i[bad[:5]], ts[bad[:5]], want_t[bad[:5]], rays[bad[:5]])
        assert np.all(ts[noisy] >= want_t[noisy])
    b.close(); s.close()


def test_degenerate_triangle_noise_is_what_it_is_documented_to_be(lib, oracle, emul):
    """the case the property test above found (and leaves out): a collinear triangle at coordinates ~100 reports a 'hit' at t = 0.25
    far outside its box (|a| = 1e-3 of rounding noise against the reference's 1e-5 threshold, modelLoader.h:64).  Pinned here so that
    the documented limitation stays a fact: the brute force over all triangles reports it, a traversal — the wide
    one and the reference's own — reports it only when the ray passes the boxes above the triangle."""
    v = np.array([[-41.84124, 84.14179, -102.99272], [-86.58763, 100.6843, 19.450985], [-64.21443, 92.41305, -41.770866],   # collinear up to rounding
                  [-50, -50, 200], [50, -50, 200], [0, 50, 200]], np.float32)                                                # an honest one
    idx = np.array([[0, 1, 2], [3, 4, 5]], np.int32)
    verts = np.ascontiguousarray(v).view(api.VEC3).reshape(-1)
    tris = np.zeros(2, dtype=api.TRIANGLE)
    tris["v0"], tris["v1"], tris["v2"] = idx[:, 0], idx[:, 1], idx[:, 2]
    mats = np.zeros(1, dtype=api.MATERIAL)
    s = api.SceneHandle(); s.set_arrays(verts, tris, mats, np.zeros(0, np.int32), 0.0)
    sv, stt, sm, sl = s.arrays()
    osc = oracle.OracleScene(sv, stt, sm, sl, 0.0, np.zeros(1, api.BVH2_NODE))
    rng = np.random.default_rng(3)
    o = rng.uniform(-600, 600, size=(20000, 3)); d = rng.normal(size=(20000, 3)) * 400
    rays = np.concatenate([o, d], 1).astype(np.float32)
    rays[0] = (89.57445, -585.1722, 78.32008, -164.97548, 681.7209, -89.48003)      # the ray hypothesis found: 'hit' at t = 0.25
    noisy = osc.hits_outside_box(rays)
    bi, bt = osc.brute_force(rays, None)
    b = s.build_bvh()
    nodes, prims = b.arrays()
    ids, ts, _ = emul_trace(emul, nodes, prims, b.info()["depth"], rays, fast=True)
    same = (ids == bi) & (ts.view(np.uint32) == bt.view(np.uint32))
    assert np.all(same[~noisy])                      # every ray without noise: bit for bit
    assert np.all(ts[noisy] >= bt[noisy])            # with noise: never a nearer hit than exists
    assert noisy[0] and bi[0] == 0 and noisy.sum() >= 1
    print(f"rays with a hit outside its triangle's box: {noisy.sum()} of {len(rays)}, of which the wide traversal differs on {(~same & noisy).sum()}")


def test_sliver_triangle_noise_close_to_the_plane(lib, oracle, emul):
    """The second case the property tests found (round 2, GPU run with 600 examples; tests/golden/sliver_scene.npz is that scene):
    triangle 17 has vertices (-2, -4, 2), (1, 1, 2) and (-0.5000298, -1.4998645, 2.0000086) — the third one 3e-5 off the segment
    between the other two — and on ray 170 triIntersect reports t = 0.99888 where the ray meets the triangle's plane at t = 0.999998:
    a point 0.0033 in front of the triangle's box (which is 8.6e-6 thick on that axis).  The brute force takes it (it is nearer
    than triangle 15's honest hit at 0.99983); a tree culls the sliver's box, whose entry distance lies behind that hit.  The first
    version of orc_hits_outside_box allowed 1e-3 of the triangle's largest coordinate (0.004 here) and called this ray clean; the
    margin is now what rounding explains (4.8e-7 of the magnitudes that enter the plane distance), and the ray is noise."""
    d = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "sliver_scene.npz"))
    v, idx, seed = d["v"], d["idx"], int(d["seed"])
    verts = np.ascontiguousarray(v).view(api.VEC3).reshape(-1)
    tris = np.zeros(len(idx), dtype=api.TRIANGLE)
    tris["v0"], tris["v1"], tris["v2"] = idx[:, 0], idx[:, 1], idx[:, 2]
    mats = np.zeros(1, dtype=api.MATERIAL)
    s = api.SceneHandle(); s.set_arrays(verts, tris, mats, np.zeros(0, np.int32), 0.0)
    sv, stt, sm, sl = s.arrays()
    osc = oracle.OracleScene(sv, stt, sm, sl, 0.0, np.zeros(1, api.BVH2_NODE))
    rays = make_rays(v, idx, seed)
    bi, bt = osc.brute_force(rays, None)
    noisy = osc.hits_outside_box(rays)
    assert bi[170] == 17 and abs(float(bt[170]) - 0.9988839) < 1e-6 and noisy[170]
    assert not osc.hits_outside_box(rays, 2e-3)[170]            # (0.0033 outside on an axis whose magnitudes are 1 .. 3)
    for max_leaf in (1, 3):
        b = s.build_bvh(max_leaf_tris=max_leaf)
        nodes, prims = b.arrays()
        for fast in (False, True):
            ids, ts, _ = emul_trace(emul, nodes, prims, b.info()["depth"], rays, fast=fast)
            same = (ids == bi) & (ts.view(np.uint32) == bt.view(np.uint32))
            assert np.all(same[~noisy]) and np.all(ts[noisy] >= bt[noisy])
        b.close()
    s.close()
