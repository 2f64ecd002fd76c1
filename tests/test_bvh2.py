"""cupt_bvh2_build == buildBVH() (BVH.h:443-474) byte for byte; leaf order = trace()'s visiting order."""
import os

import numpy as np
import pytest

from cudapathtracer_b200 import api, scenes
from conftest import CASES
from test_scene_ingest import load_via_obj, GOLD


@pytest.mark.parametrize("name", CASES)
def test_bvh2_matches_golden(lib, tmp_path, name):
    from conftest import build_case
    g = np.load(os.path.join(GOLD, name + ".npz"))
    s = load_via_obj(build_case(name)[0], tmp_path)
    nodes, depth = s.build_bvh2()
    assert depth == int(g["depth"]) and len(nodes) == len(g["tris"]) - 1
    assert nodes.tobytes() == g["nodes"].tobytes()


def test_bvh2_matches_reference_on_a_large_scene(lib, ref):
    sc = scenes.atrium(60000)
    s = api.SceneHandle()
    s.add_scene(sc)
    v, t, m, l = s.arrays()
    ref.reset()
    ref.set_arrays(v, t, m, l, s.total_light_area)
    rn, rd = ref.build_bvh()
    n, d = s.build_bvh2()
    assert d == rd and n.tobytes() == rn.tobytes()


def test_bvh2_duplicate_geometry_uses_positional_halving(lib, ref):
    """identical boxes put every centroid in one cell: the reference halves the list by position (BVH.h:263-288)"""
    sc = scenes.cornell_box()
    s = api.SceneHandle()
    for _ in range(5):
        s.add_scene(sc)
    v, t, m, l = s.arrays()
    ref.reset()
    ref.set_arrays(v, t, m, l, s.total_light_area)
    rn, rd = ref.build_bvh()
    n, d = s.build_bvh2()
    assert d == rd and n.tobytes() == rn.tobytes()


@pytest.mark.parametrize("name", CASES)
def test_bvh2_invariants_and_leaf_order(get_case, name):
    c = get_case(name)
    nodes, n = c.nodes2, len(c.tris)
    leaf = api.LEAF_FLAG
    seen = np.zeros(n, dtype=np.int32)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    order = []
    stack = [0]
    while stack:
        e = stack.pop()
        if e & leaf:
            seen[e ^ leaf] += 1
            order.append(e ^ leaf)
        else:
            stack.append(int(nodes["right"][e]))
            stack.append(int(nodes["left"][e]))
    assert np.all(seen == 1)                     # every triangle reachable exactly once
    assert len(nodes) == n - 1
    rank = c.rank
    assert np.array_equal(rank[np.array(order)], np.arange(n, dtype=np.uint32))
    # node boxes contain their triangles' vertices (children are contained transitively)
    for e in range(min(len(nodes), 2000)):
        for link in (int(nodes["left"][e]), int(nodes["right"][e])):
            if link & leaf:
                t = c.tris[link ^ leaf]
                p = pos[[t["v0"], t["v1"], t["v2"]]]
                assert np.all(p >= nodes["lo"][e] - 0) and np.all(p <= nodes["hi"][e] + 0)


def test_depth_limit_is_an_error_code(lib):
    import ctypes as C
    nodes = np.zeros(1, dtype=api.BVH2_NODE)
    rank = np.zeros(2, dtype=np.uint32)
    nodes["left"], nodes["right"] = 0 | api.LEAF_FLAG, 7 | api.LEAF_FLAG   # leaf 7 outside a 2-triangle scene
    rc = lib.cupt_bvh2_leaf_order(nodes.ctypes.data_as(C.c_void_p), 1, 2, rank.ctypes.data_as(C.c_void_p))
    assert rc == -1
