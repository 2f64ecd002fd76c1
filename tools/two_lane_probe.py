#!/usr/bin/env python
"""Experiment (GPU box): does a frame rendered as TWO concurrent half-passes (two contexts = two streams and two sets of
path buffers on the same GPU) finish sooner than one pass?  The extend kernel is issue-bound, the shade kernel is
latency-bound; co-running them hides the shade time if the extend grids leave room (CUPT_EXTEND_BLOCKS_PER_SM)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from cudapathtracer_b200 import api

sc, s, cam = bench.build_scene()
bvh = s.build_bvh()
SPP = bench.SPP
ctxs = [api.Context(0) for _ in range(int(os.environ.get("LANES", "2")))]
for c in ctxs:
    c.set_scene(s); c.set_bvh(bvh); c.set_camera(cam)
    c.set_options(max_depth=bench.DEPTH, seed=bench.SEED, time_stages=0)
split = int(os.environ.get("SPLIT", "1"))   # passes per lane
def frame():
    per = SPP // (len(ctxs) * split)
    k = 0
    for _ in range(split):
        for c in ctxs:
            c.render(k * per, per); k += 1
    for c in ctxs:
        c.sync()
for c in ctxs: c.reset_accum()
frame()
best = 1e9
for _ in range(5):
    for c in ctxs: c.reset_accum(); c.reset_stats()
    t0 = time.perf_counter(); frame(); dt = time.perf_counter() - t0
    best = min(best, dt)
rays = sum(c.stats()["rays"] for c in ctxs)
img = sum(c.read_accum(False).astype(np.float64) for c in ctxs)
print(f"lanes {len(ctxs)} split {split} blocks/SM {os.environ.get('CUPT_EXTEND_BLOCKS_PER_SM', 'max')}: {best * 1e3:.2f} ms  {rays / best / 1e6:.0f} Mrays/s  image sum {img.sum():.6e}", flush=True)
