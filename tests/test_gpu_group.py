"""cupt_group (include/cupt.h "several GPUs"): one context and one host thread per GPU, scene / BVH broadcast from rank 0 over
NCCL, sample ranges split over the members, one ncclReduce of the f32 accumulators onto rank 0.

The reference is single-device (kernel.cu:569), so the checker is the single-GPU path of this library, itself pinned to the
oracle by the other GPU tests: the group's image must equal ONE context rendering the same sample ranges one after the other
(bit for bit: the reduce adds the same partial sums), the replicated scene / BVH must give the same hits on every member,
and a progressive render must keep its running total on rank 0.  Tests that need two GPUs skip below two
(`gpurun --gpus 2 -- python -m pytest tests/test_gpu_group.py -m gpu`)."""
import ctypes as C
import multiprocessing as mp
import os
import subprocess

import numpy as np
import pytest

from cudapathtracer_b200 import api, scenes, sharding

pytestmark = pytest.mark.gpu


def device_count():
    rt = C.CDLL("libcudart.so.12")
    n = C.c_int(0)
    return n.value if rt.cudaGetDeviceCount(C.byref(n)) == 0 else 0


def case_objects(get_case, name="cornell_lights"):
    c = get_case(name)
    if not hasattr(c, "bvh8"):
        c.bvh8 = c.handle.build_bvh(tie_rank=c.rank)
    return c


def single_gpu_reference(c, ranges, **opts):
    ctx = api.Context(0)
    try:
        ctx.set_scene(c.handle); ctx.set_bvh(c.bvh8); ctx.set_bvh2(c.nodes2, c.depth2); ctx.set_camera(c.cam)
        ctx.set_options(**opts)
        hits = ctx.primary_hits()
        ctx.reset_accum()
        for first, n in ranges:
            ctx.render(first, n)
        return ctx.read_accum(False).copy(), hits, ctx.read_image_rgb8()
    finally:
        ctx.close()


@pytest.mark.parametrize("mode", [api.TRAVERSAL_WIDE, api.TRAVERSAL_WIDE_STRICT])
def test_group_of_one_equals_a_plain_context(lib, get_case, mode):
    c = case_objects(get_case)
    opts = dict(max_depth=5, seed=1234, traversal=mode)
    want, hits, rgb = single_gpu_reference(c, [(0, 5)], **opts)
    g = api.Group(devices=[0])
    try:
        assert g.info() == {"nranks": 1, "nlocal": 1, "first_rank": 0}
        g.set_scene(c.handle); g.set_bvh(c.bvh8); g.set_bvh2(c.nodes2, c.depth2); g.set_camera(c.cam)
        g.set_options(**opts)
        g.reset_accum(); g.render(0, 5); g.sync()
        assert g.read_accum(False).tobytes() == want.tobytes()
        assert np.array_equal(g.read_image_rgb8(), rgb)
        m0 = g.context(0)
        m0.set_options(**opts)
        got = m0.primary_hits()
        assert np.array_equal(got[0], hits[0]) and got[1].tobytes() == hits[1].tobytes()
        assert g.stats()["rays"] > 0
        # the tree built on rank 0's GPU instead of uploaded
        g.build_bvh_device(tie_rank=c.rank)
        g.reset_accum(); g.render(0, 5)
        assert g.read_accum(False).tobytes() == want.tobytes()
    finally:
        g.close()


@pytest.mark.parametrize("mode", [api.TRAVERSAL_WIDE, api.TRAVERSAL_WIDE_STRICT])
def test_group_of_all_gpus_in_one_process(lib, get_case, mode):
    n = device_count()
    if n < 2:
        pytest.skip("needs two GPUs")
    c = case_objects(get_case, "atrium20k")
    opts = dict(max_depth=5, seed=99, traversal=mode)
    spp = 2 * n + 1                                                          # uneven split
    ranges = [sharding.sample_range(r, n, spp) for r in range(n)]
    want, hits, _ = single_gpu_reference(c, ranges, **opts)
    g = api.Group(devices=list(range(n)))
    try:
        g.set_scene(c.handle); g.set_bvh(c.bvh8); g.set_bvh2(c.nodes2, c.depth2); g.set_camera(c.cam)
        g.set_options(**opts)
        for i in range(n):                                                   # the broadcast copies trace like the original
            m = g.context(i)
            m.set_options(**opts)
            got = m.primary_hits()
            assert np.array_equal(got[0], hits[0]) and got[1].tobytes() == hits[1].tobytes(), i
        g.reset_accum(); g.render(0, spp); g.sync()
        img = g.read_accum(False)
        if n == 2:
            assert img.tobytes() == want.tobytes()                           # a + b either way
        assert np.allclose(img, want, rtol=1e-6, atol=1e-6)
        # progressive: the running total stays on rank 0, the other members start from zero again
        g.render(spp, n); g.sync()
        more, _, _ = single_gpu_reference(c, ranges + [(spp + r, 1) for r in range(n)], **opts)
        assert np.allclose(g.read_accum(False), more, rtol=1e-6, atol=1e-6)
        assert np.allclose(g.read_accum(True), more / (spp + n), rtol=1e-6, atol=1e-6)
        for i in range(1, n):
            assert not g.context(i).read_accum(False).any()
    finally:
        g.close()


def _rank_process(rank, nranks, uid, conn):
    """one process per GPU (the torchrun shape): this process's member of the group"""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from cudapathtracer_b200 import api as a, scenes as sc_
    try:
        g = a.Group(device=rank, rank=rank, nranks=nranks, unique_id=uid)
        scene = sc_.cornell_box(extra_lights=8)
        cam = a.make_camera(48, 40, **scene.camera)
        if rank == 0:
            s = a.SceneHandle(); s.add_scene(scene)
            n2, d2 = s.build_bvh2()
            b = s.build_bvh(tie_rank=a.bvh2_leaf_order(n2, s.counts()[1]))
            g.set_scene(s); g.set_bvh(b)
        else:
            g.set_scene(None); g.set_bvh(None)                               # arrives by broadcast
        g.set_camera(cam)
        g.set_options(max_depth=5, seed=7)
        g.reset_accum(); g.render(0, 6); g.sync()
        img = g.read_accum(False)
        assert (img is None) == (rank != 0)
        conn.send(("ok", None if img is None else img.tobytes()))
        g.close()
    except Exception as e:   # noqa: BLE001
        conn.send(("error", repr(e)))


def test_group_one_process_per_gpu(lib, get_case):
    n = device_count()
    if n < 2:
        pytest.skip("needs two GPUs")
    n = 2
    uid = api.group_unique_id()
    ctx = mp.get_context("spawn")
    pipes, procs = [], []
    for r in range(n):
        a, b = ctx.Pipe()
        p = ctx.Process(target=_rank_process, args=(r, n, uid, b))
        p.start(); pipes.append(a); procs.append(p)
    results = []
    for a, p in zip(pipes, procs):
        assert a.poll(180), "a rank did not answer"
        results.append(a.recv())
        p.join(60)
    assert all(r[0] == "ok" for r in results), results
    c = case_objects(get_case, "cornell_lights")
    want, _, _ = single_gpu_reference(c, [sharding.sample_range(r, n, 6) for r in range(n)], max_depth=5, seed=7)
    assert results[0][1] == want.tobytes() and results[1][1] is None


def test_render_loop_example_on_all_gpus(lib, tmp_path):
    """examples/render_loop.cpp with its gpus argument: a C++ main() written like the reference's reaches every GPU without
    Python; its PPM is byte-identical to the single-context render of the same sample ranges"""
    n = max(1, min(device_count(), 8))
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "examples", "render_loop")
    sc = scenes.cornell_box()
    obj, mtl = scenes.to_obj(sc)
    (tmp_path / "models").mkdir()
    (tmp_path / "models" / "cornell.obj").write_text(obj)
    (tmp_path / "models" / "scene.mtl").write_text(mtl)
    out = tmp_path / "group.ppm"
    spp = 2 * n
    r = subprocess.run([exe, str(tmp_path / "models" / "cornell.obj"), str(tmp_path / "models") + "/", "64", "48", str(spp), "5", str(out), "wide", str(n)],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert r.returncode == 0, r.stdout
    assert f"{n} GPUs" in r.stdout and "Mrays/s (counted)" in r.stdout
    s = api.SceneHandle()
    s.load_obj(str(tmp_path / "models" / "cornell.obj"), str(tmp_path / "models") + "/")
    n2, d2 = s.build_bvh2()
    b = s.build_bvh(tie_rank=api.bvh2_leaf_order(n2, s.counts()[1]))
    ctx = api.Context(0)
    try:
        ctx.set_scene(s); ctx.set_bvh(b); ctx.set_camera(api.make_camera(64, 48, **sc.camera))
        ctx.set_options(traversal=api.TRAVERSAL_WIDE, max_depth=5, seed=1234)
        ctx.reset_accum()
        for rr in range(n):
            ctx.render(*sharding.sample_range(rr, n, spp))
        ref = tmp_path / "ref.ppm"
        api.write_ppm(str(ref), ctx.read_accum(True), 64, 48)
    finally:
        ctx.close()
    if n <= 2:
        assert out.read_bytes() == ref.read_bytes()
    else:   # more than two partial sums: the reduce may add them in another order (1e-7 relative), a level may flip on a boundary
        a, bb = np.array(out.read_text().split()[4:], int), np.array(ref.read_text().split()[4:], int)
        assert len(a) == len(bb) and np.abs(a - bb).max() <= 1 and (a != bb).mean() <= 1e-3
