import sys, os
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
from cudapathtracer_b200 import api, scenes
import oraclebind
sc = scenes.atrium()
s = api.SceneHandle(); s.add_scene(sc)
n2, d2 = s.build_bvh2()
v,t,m,l = s.arrays()
rank = api.bvh2_leaf_order(n2, len(t))
b = s.build_bvh(tie_rank=rank)
cam = api.make_camera(1024, 1024, **sc.camera)
ctx = api.Context(0)
ctx.set_scene(s); ctx.set_bvh(b); ctx.set_bvh2(n2, d2); ctx.set_camera(cam)
ctx.set_options(traversal=api.TRAVERSAL_WIDE, max_depth=5, seed=1234)
wi, wt = ctx.primary_hits()
ctx.set_options(traversal=api.TRAVERSAL_REF_BVH2)
ri, rt = ctx.primary_hits()
print("wide misses", (wi<0).sum(), "wt>=60", (wt>=60).sum(), "ref misses", (ri<0).sum(), "rt>=60", (rt>=60).sum())
bad = np.nonzero((wi<0)|(wt>=60))[0]
print(bad[:20], [(int(i)%1024, int(i)//1024) for i in bad[:10]])
rays = ctx.camera_rays()
osc = oraclebind.OracleScene(v,t,m,l,s.total_light_area,n2)
if len(bad):
    bi, bt = osc.brute_force(rays[bad[:50]], rank)
    print("brute", bi[:10], bt[:10], "wide", wi[bad[:10]], wt[bad[:10]])
