"""Developer tool (GPU box): run the legacy kernel.cu baseline on config 2 and compare its image with ours."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from cudapathtracer_b200 import api, scenes
import legacybind
sc = scenes.atrium(262144)
s = api.SceneHandle(); s.add_scene(sc)
cam = api.make_camera(1024, 1024, **sc.camera)
v, t, m, l = s.arrays()
for spp in (4, 16):
    r = legacybind.render(v, t, m, l, s.total_light_area, cam, spp, want_image=True)
    print(f"legacy kernel.cu sm_100a: {spp} spp in {r['ms']:.1f} ms, rays {r['rays']} ({r['rays']/(1024*1024*spp):.2f}/sample) -> "
          f"{r['rays']/r['ms']/1e3:.1f} Mrays/s, {1024*1024*spp/r['ms']/1e3:.1f} Msamples/s, bvh depth {r['depth']} build {r['build_s']:.2f}s, mean {r['image'].mean():.4f}", flush=True)
