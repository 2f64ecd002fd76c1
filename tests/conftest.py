import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def lib():
    """libcupt.so must exist: there is no fallback to fall back to"""
    from cudapathtracer_b200 import api
    if not os.path.exists(api.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return api.load_library()


@pytest.fixture(scope="session")
def oracle():
    import oraclebind
    if not oraclebind.available():
        import __graft_entry__
        __graft_entry__.build()
    return oraclebind


@pytest.fixture(scope="session")
def ref():
    """the real reference compiled for the host (only where /root/reference was available at build time)"""
    import refbind
    if not refbind.available():
        pytest.skip("oracle/_ref/libcupt_ref.so not built (no /root/reference on this machine)")
    return refbind


def build_case(name):
    """(scenes.Scene, width, height) of the small named parity cases shared by golden/CPU/GPU tests"""
    from cudapathtracer_b200 import scenes
    if name == "cornell":
        return scenes.cornell_box(), 64, 64
    if name == "cornell_lights":
        return scenes.cornell_box(extra_lights=8), 48, 40
    if name == "atrium20k":
        return scenes.atrium(20000), 64, 64
    if name == "soup5k":
        return scenes.triangle_soup(5000, extent=2.0, edge=(0.1, 0.6), cam_z=7.0), 64, 48
    raise KeyError(name)


CASES = ["cornell", "cornell_lights", "atrium20k", "soup5k"]


class Case:
    """host-side objects of one parity case: product scene handle, reference-layout BVH2, ranks, oracle scene"""

    def __init__(self, name, oraclebind):
        from cudapathtracer_b200 import api
        self.name = name
        self.scene, self.w, self.h = build_case(name)
        self.handle = api.SceneHandle()
        self.handle.add_scene(self.scene)
        self.verts, self.tris, self.mats, self.lights = self.handle.arrays()
        self.area = self.handle.total_light_area
        self.nodes2, self.depth2 = self.handle.build_bvh2()
        self.rank = api.bvh2_leaf_order(self.nodes2, len(self.tris))
        self.cam = api.make_camera(self.w, self.h, **self.scene.camera)
        self.osc = oraclebind.OracleScene(self.verts, self.tris, self.mats, self.lights, self.area, self.nodes2)


_case_cache = {}


@pytest.fixture(scope="session")
def get_case(lib, oracle):
    def get(name):
        if name not in _case_cache:
            _case_cache[name] = Case(name, oracle)
        return _case_cache[name]
    return get
