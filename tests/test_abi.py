"""C-ABI surface: the library loads, exports every symbol include/cupt.h declares, PODs have the reference's
sizes.  No compute calls (the CPU suite has no GPU)."""
import ctypes as C
import os
import re
import struct

import numpy as np
import pytest

from cudapathtracer_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "cupt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(cupt_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol(lib):
    syms = declared_symbols()
    assert len(syms) >= 40
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_header_cites_reference_interfaces():
    text = open(os.path.join(ROOT, "include", "cupt.h")).read()
    for cite in ("modelLoader.h:125", "BVH.h:443", "kernel.cu:709-736", "camera.h:26-35", "kernel.cu:763-778"):
        assert cite in text


def test_pod_sizes_match_reference_layout(lib):
    out = (C.c_int * 7)()
    assert lib.cupt_struct_sizes(out) == 0
    assert list(out)[:6] == [12, 28, 48, 32, 32, 64]   # SURVEY.md appendix A (sizeof probe of the reference headers)
    assert lib.cupt_version() == 100


def test_pod_sizes_against_compiled_reference(ref):
    s = ref.struct_sizes()
    assert (s["vec3"], s["triangle"], s["materialDesc"], s["camera"], s["BVH_array_node"]) == (12, 28, 48, 32, 32)
    assert (api.VEC3.itemsize, api.TRIANGLE.itemsize, api.MATERIAL.itemsize, api.CAMERA.itemsize, api.BVH2_NODE.itemsize) == (12, 28, 48, 32, 32)


def test_parallel_epsilon_constant():
    """`abs(a) < 0.00001` (modelLoader.h:64) compares a float with a double literal; traverse.cuh uses the float
    threshold CUPT_PARALLEL_EPS_BITS = smallest float >= 0.00001"""
    text = open(os.path.join(ROOT, "cudapathtracer_b200", "csrc", "traverse.cuh")).read()
    bits = int(re.search(r"CUPT_PARALLEL_EPS_BITS (0x[0-9A-Fa-f]+)u", text).group(1), 16)
    thr = struct.unpack("<f", struct.pack("<I", bits))[0]
    below = struct.unpack("<f", struct.pack("<I", bits - 1))[0]
    assert float(thr) >= 0.00001 and float(below) < 0.00001
    # so for every float f: (double)f < 0.00001  <=>  f < thr
    for f in (below, thr, np.float32(1e-5), np.nextafter(np.float32(1e-5), np.float32(1))):
        assert (float(f) < 0.00001) == (np.float32(f) < np.float32(thr))


def test_errors_are_codes_not_exits(lib):
    assert lib.cupt_scene_counts(None, None) == -1
    assert b"bad arguments" in lib.cupt_last_error()
    s = api.SceneHandle()
    with pytest.raises(api.CuptError, match="Cannot open file"):
        s.load_obj("/nonexistent/file.obj")
    with pytest.raises(api.CuptError, match="at least 2 triangles"):
        s.build_bvh2()
    with pytest.raises(api.CuptError, match="no primitives"):
        s.build_bvh()


def test_no_cpu_fallback(lib):
    """without a CUDA device the context refuses to exist (on a GPU box it is created and destroyed)"""
    h = C.c_void_p()
    rc = lib.cupt_create(0, C.byref(h))
    if rc == 0:
        lib.cupt_destroy(h)
    else:
        assert rc == -3 and b"no CPU fallback" in lib.cupt_last_error()


def test_product_never_touches_the_oracle():
    """nothing under cudapathtracer_b200/ or include/ may import, link or open oracle/"""
    bad = []
    for base in ("cudapathtracer_b200", "include"):
        for dp, _, fs in os.walk(os.path.join(ROOT, base)):
            if "build" in dp.split(os.sep):
                continue
            for f in fs:
                if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h", "Makefile")):
                    t = open(os.path.join(dp, f), errors="replace").read()
                    if re.search(r"liboracle|oraclebind|refbind|libcupt_ref|oracle/oracle\.c", t):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_compat_shim_compiles_against_the_reference_headers(lib):
    """include/cupt_compat.h + examples/reference_main.cpp (loadOBJ, buildBVH, the camera POD, MAX_BVH_DEPTH, the reference's
    color functions in the PPM writer) compile with a plain host compiler against the reference's OWN vec3.h / color.h /
    camera.h.  Built by __graft_entry__.build() where /root/reference exists; run on the GPU by test_gpu_parity."""
    import os
    if not os.path.isdir("/root/reference"):
        pytest.skip("no reference tree on this machine")
    import __graft_entry__
    __graft_entry__.build()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    assert os.path.exists(os.path.join(root, "examples", "reference_main"))
    text = open(os.path.join(root, "include", "cupt_compat.h")).read()
    for name in ("loadOBJ", "buildBVH", "BVH_array", "materialDesc", "totalLightArea"):
        assert name in text
