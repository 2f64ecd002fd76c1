"""Structural invariants of the wide BVH image (Node8 / Prim48 arrays), shared by the host-builder tests (CPU) and the
device-builder tests (GPU)."""
import numpy as np


def decode_boxes(nodes):
    scale = np.ldexp(1.0, nodes["e"].astype(np.int64) - 127)                      # [n,3]
    lo = nodes["p"][:, :, None].astype(np.float64) + nodes["qlo"].astype(np.float64) * scale[:, :, None]
    hi = nodes["p"][:, :, None].astype(np.float64) + nodes["qhi"].astype(np.float64) * scale[:, :, None]
    return lo, hi                                                                   # [n,3,8]


def check_wide_bvh(c, nodes, prims, info, max_leaf):
    n = len(c.tris)
    assert info["node_bytes"] == 80 and info["tri_bytes"] == 48
    assert len(prims) == n and np.array_equal(np.sort(prims["id"]), np.arange(n, dtype=np.uint32))   # each triangle once
    assert np.array_equal(prims["rank"], c.rank[prims["id"]]) and np.all(prims["kind"] == 0)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    t = c.tris[prims["id"]]
    v0 = pos[t["v0"]]
    assert np.array_equal(prims["a"], v0)
    assert np.array_equal(prims["b"], pos[t["v1"]] - v0) and np.array_equal(prims["c"], pos[t["v2"]] - v0)   # e1, e2 as f32 subtractions
    lo, hi = decode_boxes(nodes)
    imask = nodes["imask"].astype(np.uint32)
    pvalid = nodes["pvalid"].astype(np.uint64)
    slots = np.arange(8)
    inner = ((imask[:, None] >> slots) & 1).astype(bool)                          # [n,8]
    nib = ((pvalid[:, None] >> (4 * slots).astype(np.uint64)) & np.uint64(0xF)).astype(np.int64)   # valid-record nibble per slot
    assert np.all(np.isin(nib, (0, 1, 3, 7))) and np.all(nib[inner] == 0) and np.all(nodes["reserved"] == 0)
    # walk: inner children consecutive in slot order, every node reachable once, boxes nested and conservative
    seen = np.zeros(len(nodes), np.int32)
    tri_seen = np.zeros(n, np.int32)
    stack = [(0, [], 1)]
    maxd = 0
    vmin = np.minimum(np.minimum(pos[t["v0"]], pos[t["v1"]]), pos[t["v2"]]).astype(np.float64)
    vmax = np.maximum(np.maximum(pos[t["v0"]], pos[t["v1"]]), pos[t["v2"]]).astype(np.float64)
    while stack:
        i, anc, d = stack.pop()
        seen[i] += 1
        maxd = max(maxd, d)
        k = 0
        ofs = 0
        for s in range(8):
            if not inner[i, s] and nib[i, s] == 0:
                continue
            blo, bhi = lo[i, :, s], hi[i, :, s]
            assert np.all(blo <= bhi)
            if inner[i, s]:
                stack.append((int(nodes["child_base"][i]) + k, anc + [(blo, bhi)], d + 1))
                k += 1
            else:
                cnt = {1: 1, 3: 2, 7: 3}[int(nib[i, s])]
                assert cnt <= max_leaf
                first = int(nodes["tri_base"][i]) + ofs                             # records are consecutive in slot order
                ofs += cnt
                for r in range(first, first + cnt):
                    tri_seen[r] += 1
                    for alo, ahi in anc + [(blo, bhi)]:                            # every ancestor's decoded box is conservative
                        assert np.all(vmin[r] >= alo) and np.all(vmax[r] <= ahi)
    assert np.all(seen == 1) and np.all(tri_seen == 1) and maxd == info["depth"]
