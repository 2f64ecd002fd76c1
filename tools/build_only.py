"""Developer tool (GPU box): build the wide BVH of an N-million-triangle soup on the device twice (first = warm-up) and print the
stage times; meant for an ncu launch list of the builder's kernels:  SIZE=30 python tools/build_only.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cudapathtracer_b200 import api, scenes

n = int(float(os.environ.get("SIZE", "30")) * 1e6)
sc = scenes.triangle_soup(n) if os.environ.get("SCENE", "soup") == "soup" else scenes.atrium(n)
s = api.SceneHandle(); s.add_scene(sc)
ctx = api.Context(0)
ctx.set_scene(s)
if os.environ.get("L2FETCH"):   # experiment: cudaLimitMaxL2FetchGranularity (0x05), a hint of 32 / 64 / 128 bytes
    import ctypes as C
    rt = C.CDLL("libcudart.so.12")
    print("cudaDeviceSetLimit ->", rt.cudaDeviceSetLimit(5, C.c_size_t(int(os.environ["L2FETCH"]))))
    v = C.c_size_t(0); rt.cudaDeviceGetLimit(C.byref(v), 5); print("L2 fetch granularity now", v.value)
for _ in range(int(os.environ.get("REPS", "2"))):
    info = ctx.build_bvh_device(radius=int(os.environ.get("RADIUS", "0")))
print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in info.items()})
if os.environ.get("HASH") and n <= 40_000_000:   # the tree's bytes (nodes + records), to check that a builder change leaves the trees as they were
    import hashlib
    nodes, prims, *_ = ctx.read_bvh()
    print("tree sha1", hashlib.sha1(nodes.tobytes()).hexdigest()[:16], hashlib.sha1(prims.tobytes()).hexdigest()[:16])
