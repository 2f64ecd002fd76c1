"""Wide BVH builder (host) — structural invariants, and the traversal header checked on the CPU through the
emulation harness tests/emul (structural check of the host-built data; the kernels themselves are GPU tests)."""
import ctypes as C
import os

import numpy as np
import pytest

from cudapathtracer_b200 import api, scenes
from conftest import CASES

EMUL = os.path.join(os.path.dirname(os.path.abspath(__file__)), "emul", "libemul.so")


def P(a):
    return a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def emul(lib):
    if not os.path.exists(EMUL):
        import __graft_entry__
        __graft_entry__.build()
    return C.CDLL(EMUL)


def emul_trace(emul, nodes, prims, depth, rays, fast=False):
    """fast: through wide_node_step_fast and the sparse primitive groups (what k_extend runs) instead of the definition"""
    rays = np.ascontiguousarray(rays, dtype=np.float32)
    ids, ts = np.zeros(len(rays), np.int32), np.zeros(len(rays), np.float32)
    cnt = (C.c_ulonglong * 2)()
    fn = emul.emul_trace_wide_fast if fast else emul.emul_trace_wide
    fn(P(nodes), P(prims), depth, P(rays), C.c_longlong(len(rays)), P(ids), P(ts), cnt)
    return ids, ts, (cnt[0], cnt[1])


def random_rays(c, n, seed):
    rng = np.random.default_rng(seed)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    o = rng.uniform(pos.min(0), pos.max(0), size=(n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d], axis=1).astype(np.float32)


from bvhcheck import check_wide_bvh, decode_boxes  # noqa: E402


@pytest.mark.parametrize("name", CASES)
@pytest.mark.parametrize("max_leaf", [1, 3])
def test_bvh8_invariants(get_case, name, max_leaf):
    c = get_case(name)
    b = c.handle.build_bvh(max_leaf_tris=max_leaf, tie_rank=c.rank)
    info = b.info()
    nodes, prims = b.arrays()
    check_wide_bvh(c, nodes, prims, info, max_leaf)


@pytest.mark.parametrize("name", CASES)
def test_wide_traversal_logic_matches_brute_force_and_reference(get_case, emul, name):
    c = get_case(name)
    b = c.handle.build_bvh(tie_rank=c.rank)
    nodes, prims = b.arrays()
    rays = c.osc.camera_rays(c.cam)
    rng = np.random.default_rng(11)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    o = rng.uniform(pos.min(0), pos.max(0), size=(4000, 3))
    d = rng.normal(size=(4000, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    axis = np.zeros((64, 6)); axis[:, :3] = rng.uniform(pos.min(0), pos.max(0), size=(64, 3)); axis[np.arange(64), 3 + np.arange(64) % 3] = 1.0  # axis-parallel: 1/0 in the slab test
    rays = np.concatenate([rays, np.concatenate([o, d], axis=1), axis]).astype(np.float32)
    ids, ts, (nn, nt) = emul_trace(emul, nodes, prims, b.info()["depth"], rays)
    bi, bt = c.osc.brute_force(rays, c.rank)
    assert np.array_equal(ids, bi) and ts.tobytes() == bt.tobytes()               # == argmin over all triangles
    fi, ft, (fn, fnt) = emul_trace(emul, nodes, prims, b.info()["depth"], rays, fast=True)
    assert np.array_equal(fi, bi) and ft.tobytes() == bt.tobytes()                # the hot kernel's node step: same hits
    assert fn == nn                                                                # ... through the same nodes
    ri, rt = c.osc.trace_rays(rays)
    differ = (ids != ri) | (ts.view(np.uint32) != rt.view(np.uint32))
    assert differ.sum() <= 4 and np.all(ts[differ] <= rt[differ])                 # only hits the reference dropped
    assert nn > 0 and nt > 0


@pytest.mark.parametrize("name", ["atrium20k", "soup5k"])
def test_insertion_based_optimisation_keeps_the_tree_valid_and_makes_it_cheaper(get_case, emul, name):
    """cupt_bvh_opts.reinsert_passes (default 2 sweeps): subtrees are moved to where the binary tree's total surface area drops
    most (bvh8_build.cpp: Reinserter).  Whatever the sweeps do, the wide BVH must keep every structural invariant, return the
    same hits (argmin over all triangles), and the SAH cost and the node visits per ray must not get worse."""
    c = get_case(name)
    rays = random_rays(c, 6000, 5)
    want = c.osc.brute_force(rays, c.rank)
    stats = {}
    for passes in (-1, 1, 0, 6):          # none, one sweep, the default (two), many
        b = c.handle.build_bvh(tie_rank=c.rank, reinsert_passes=passes)
        info = b.info()
        nodes, prims = b.arrays()
        check_wide_bvh(c, nodes, prims, info, 3)
        ids, ts, (nn, nt) = emul_trace(emul, nodes, prims, info["depth"], rays, fast=True)
        assert np.array_equal(ids, want[0]) and ts.tobytes() == want[1].tobytes(), passes
        stats[passes] = (info["sah_cost"], nn)
    assert stats[0][0] < stats[-1][0] and stats[6][0] <= stats[1][0] * 1.001, stats       # cheaper by the builder's own measure
    assert stats[0][1] < stats[-1][1], stats                                                  # and fewer node visits on these rays


def test_bvh2_traversal_header_is_the_reference_algorithm(get_case, emul):
    c = get_case("atrium20k")
    rays = c.osc.camera_rays(c.cam)
    ids, ts = np.zeros(len(rays), np.int32), np.zeros(len(rays), np.float32)
    vf = np.ascontiguousarray(c.verts.view(np.float32))
    emul.emul_trace_bvh2(P(c.nodes2), P(vf), P(c.tris), P(rays), C.c_longlong(len(rays)), P(ids), P(ts))
    oi, ot = c.osc.trace_rays(rays)
    assert np.array_equal(ids, oi) and ts.tobytes() == ot.tobytes()


def test_tie_break_follows_reference_leaf_order(lib, oracle, emul):
    """two coincident triangles: the reference returns the one its traversal meets first (kernel.cu:127)"""
    sc = scenes.cornell_box()
    s = api.SceneHandle()
    s.add_scene(sc)
    s.add_scene(sc)                        # every triangle twice, exact ties everywhere
    v, t, m, l = s.arrays()
    n2, d2 = s.build_bvh2()
    rank = api.bvh2_leaf_order(n2, len(t))
    osc = oracle.OracleScene(v, t, m, l, s.total_light_area, n2)
    cam = api.make_camera(48, 48, **sc.camera)
    rays = osc.camera_rays(cam)
    ri, rt = osc.trace_rays(rays)
    b = s.build_bvh(tie_rank=rank)
    nodes, prims = b.arrays()
    ids, ts, _ = emul_trace(emul, nodes, prims, b.info()["depth"], rays)
    assert np.array_equal(ids, ri) and ts.tobytes() == rt.tobytes()
    b2 = s.build_bvh()                     # without ranks the lower triangle index wins instead
    nodes, prims = b2.arrays()
    ids2, ts2, _ = emul_trace(emul, nodes, prims, b2.info()["depth"], rays)
    assert ts2.tobytes() == rt.tobytes() and np.all(ids2[ids2 >= 0] < 32)


def test_large_build_is_deterministic_and_parallel(lib):
    sc = scenes.triangle_soup(200000)
    s = api.SceneHandle()
    s.add_scene(sc)
    a = s.build_bvh(threads=1).arrays()
    b = s.build_bvh(threads=8).arrays()
    assert a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()


def test_spheres_in_the_wide_bvh_match_the_oracle_restatement(get_case, emul):
    """config 1's spheres (sphere.h:7-13): PARITY UNPINNED — the reference has no intersector; the wide traversal is
    checked against the oracle's restatement of the product's own sphere arithmetic (brute force over all primitives)"""
    c = get_case("cornell_spheres")
    b = c.handle.build_bvh(tie_rank=c.rank)
    info = b.info()
    assert info["num_spheres"] == 2 and info["num_tri_records"] == len(c.tris) + 2
    nodes, prims = b.arrays()
    assert sorted(prims["id"][prims["kind"] == 1].tolist()) == [0, 1]
    rays = np.concatenate([c.osc.camera_rays(c.cam), random_rays(c, 20000, 9)])
    ids, ts, _ = emul_trace(emul, nodes, prims, info["depth"], rays)
    bi, bt = c.osc.brute_force(rays, c.rank)
    assert np.array_equal(ids, bi) and ts.tobytes() == bt.tobytes()
    assert (ids < -1).sum() > 500          # spheres are actually hit (id = 0x80000000 | index)
    oi, ot = c.osc.trace_rays(rays)         # the reference traversal + spheres agrees except for reference-dropped triangles
    assert ((oi != ids) | (ot.view(np.uint32) != ts.view(np.uint32))).sum() <= 4
