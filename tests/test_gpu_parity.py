"""GPU parity tests: the CUDA path, called through the C-ABI, against the oracle and the golden vectors.

Tolerances (also in DESIGN.md §Parity):
  * primary hit ids and raw t: BIT-EXACT in CUPT_TRAVERSAL_REF_BVH2 mode, unconditionally; in the wide mode the
    result equals argmin over all triangles (oracle brute force) bit for bit, and therefore equals the reference
    except on rays where the reference's non-conservative slab test (BVH.h:65,79) dropped the winning triangle —
    those rays are counted and must all be explained by the brute force.
  * radiance (f32 sums on the GPU vs f64 in the reference): per-pixel |gpu - oracle| <= 1e-3 * (1 + oracle) on at
    least 99% of pixels and relative RMSE <= 5% at equal spp and seed (paths are identical until a device
    sinf/cosf ulp or a reference-dropped hit changes a bounce); in the REF_BVH2 and strict modes >= 99.9% / 1%.  In
    the wide mode a ray on which the reference lost the nearest triangle continues as a different path; the strict mode
    lists exactly those rays (cupt_read_retrace_counts), the pixels they belong to are masked and the rest is held to
    the same 99.9% / 1%; the whole image's mean stays within 3%.
"""
import os

import numpy as np
import pytest

from cudapathtracer_b200 import api, scenes
from conftest import CASES

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture(scope="module")
def ctx(lib):
    c = api.Context(0)
    yield c
    c.close()


def load(ctx, case, mode, **opts):
    ctx.set_scene(case.handle)
    if not hasattr(case, "bvh8"):
        case.bvh8 = case.handle.build_bvh(tie_rank=case.rank)
    ctx.set_bvh(case.bvh8)
    ctx.set_bvh2(case.nodes2, case.depth2)
    ctx.set_camera(case.cam)
    ctx.set_options(traversal=mode, **opts)


def hits_equal(a, b):
    return np.array_equal(a[0], b[0]) and a[1].tobytes() == b[1].tobytes()


@pytest.mark.parametrize("name", CASES)
def test_camera_rays_bit_exact(ctx, get_case, name):
    c = get_case(name)
    ctx.set_camera(c.cam)
    assert ctx.camera_rays().tobytes() == c.osc.camera_rays(c.cam).tobytes()


@pytest.mark.parametrize("name", CASES)
def test_primary_hits_bit_exact_vs_golden_and_oracle(ctx, get_case, name):
    c = get_case(name)
    g = np.load(os.path.join(GOLD, name + ".npz"))
    load(ctx, c, api.TRAVERSAL_REF_BVH2)
    ids, ts = ctx.primary_hits()
    oi, ot = c.osc.primary_hits(c.cam)
    assert np.array_equal(ids, oi) and ts.tobytes() == ot.tobytes()
    # the goldens were produced through the OBJ route (vertex renumbering does not change any triangle)
    assert np.array_equal(ts.view(np.uint32), g["primary_t_bits"])
    assert np.array_equal(ids, g["primary_ids"])
    load(ctx, c, api.TRAVERSAL_WIDE)
    wi, wt = ctx.primary_hits()
    rays = c.osc.camera_rays(c.cam)
    bi, bt = c.osc.brute_force(rays, c.rank)
    assert np.array_equal(wi, bi) and wt.tobytes() == bt.tobytes()
    differ = (wi != oi) | (wt.view(np.uint32) != ot.view(np.uint32))
    assert differ.sum() <= 2 and np.all(wt[differ] <= ot[differ])


@pytest.mark.parametrize("name", CASES)
def test_incoherent_rays(ctx, get_case, name):
    c = get_case(name)
    rng = np.random.default_rng(5)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    n = 30000
    o = rng.uniform(pos.min(0), pos.max(0), size=(n, 3))
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    rays[:90, 3:] = np.tile(np.eye(3, dtype=np.float32), (30, 1))          # axis-parallel directions (1/0 slabs)
    rays[90:96, 3:] = 0.0                                                   # degenerate direction: no hit, no hang
    oi, ot = c.osc.trace_rays(rays)
    load(ctx, c, api.TRAVERSAL_REF_BVH2)
    assert hits_equal(ctx.trace_rays(rays), (oi, ot))
    load(ctx, c, api.TRAVERSAL_WIDE)
    wi, wt = ctx.trace_rays(rays)
    ok = np.ones(n, bool); ok[90:96] = False
    bi, bt = c.osc.brute_force(rays[ok], c.rank)
    assert np.array_equal(wi[ok], bi) and wt[ok].tobytes() == bt.tobytes()
    assert np.all(wi[90:96] == -1)


def test_empty_and_ragged_inputs(ctx, get_case):
    c = get_case("cornell")
    load(ctx, c, api.TRAVERSAL_WIDE)
    ids, ts = ctx.trace_rays(np.zeros((0, 6), np.float32))
    assert len(ids) == 0
    ids, ts = ctx.trace_rays(np.array([[0, 1, 3, 0, 0, -1]], np.float32))   # one ray, not a multiple of the warp size
    one = np.array([[0, 1, 3, 0, 0, -1]], np.float32)
    assert hits_equal((ids, ts), c.osc.trace_rays(one)) and ids[0] >= 0 and 0 < ts[0] < 4.1
    ids, ts = ctx.trace_rays(np.array([[0, 1, 3, 0, 0, 1]], np.float32))    # looks away: miss = {-1, 1e5} (kernel.cu:117-118)
    assert ids[0] == -1 and ts[0] == np.float32(100000.0)
    cam = api.make_camera(37, 21, **c.scene.camera)                         # not a multiple of the 8x4 tile
    ctx.set_camera(cam)
    a = ctx.primary_hits()
    assert hits_equal(a, c.osc.primary_hits(cam))
    ctx.set_options(max_depth=1)
    ctx.reset_accum(); ctx.render(0, 0)                                      # zero samples: accumulator stays zero
    assert not ctx.read_accum(False).any()


@pytest.mark.parametrize("name", CASES)
def test_radiance_vs_oracle_and_golden(ctx, get_case, name):
    c = get_case(name)
    g = np.load(os.path.join(GOLD, name + ".npz"))
    for nb, spp, seed in ((5, 4, 1234), (3, 2, 99)):
        omean, osum, orays = c.osc.render_pt(c.cam, nb, seed, 0, spp)
        ref = osum / spp
        gold = g[f"pt_sum_nb{nb}_spp{spp}_seed{seed}"] / spp
        # pixels none of whose rays the strict mode had to re-trace: there the reference reached the nearest triangle on
        # every ray, so the plain wide mode follows the same paths and is held to the reference-exact tolerance
        load(ctx, c, api.TRAVERSAL_WIDE_STRICT, max_depth=nb, seed=seed, record_retraces=1)
        ctx.reset_accum(); ctx.render(0, spp)
        clean = ctx.read_retrace_counts() == 0
        ctx.set_options(record_retraces=0)
        assert clean.mean() >= 0.98, (name, clean.mean())
        for mode, keep in ((api.TRAVERSAL_REF_BVH2, None), (api.TRAVERSAL_WIDE_STRICT, None), (api.TRAVERSAL_WIDE, clean)):
            load(ctx, c, mode, max_depth=nb, seed=seed)
            ctx.reset_accum()
            ctx.render(0, spp)
            img = ctx.read_accum(True).astype(np.float64)
            for want in (ref, gold):
                a, w = (img, want) if keep is None else (img[keep], want[keep])
                close = np.all(np.abs(a - w) <= 1e-3 * (1 + np.abs(w)), axis=1)
                rel = np.sqrt(((a - w) ** 2).mean()) / max(w.mean(), 1e-9)
                assert close.mean() >= 0.999, (name, mode, close.mean())
                assert rel <= 0.01, (name, mode, rel)
                # the whole image, re-traced pixels included: a path that continues differently moves one pixel by emission / spp
                assert abs(img.mean() - want.mean()) <= 0.03 * want.mean(), (name, mode, img.mean(), want.mean())


def test_sample_ranges_compose(ctx, get_case):
    """rendering samples [0,6) at once == [0,2) + [2,6) == any samples_per_pass: the Philox stream is keyed by the
    sample index, which is what makes sample-range sharding across GPUs exact"""
    c = get_case("cornell_lights")
    load(ctx, c, api.TRAVERSAL_WIDE, max_depth=5, seed=7, samples_per_pass=0)
    ctx.reset_accum(); ctx.render(0, 6); a = ctx.read_accum(False)
    ctx.reset_accum(); ctx.render(0, 2); ctx.render(2, 4); b = ctx.read_accum(False)
    ctx.set_options(samples_per_pass=1)
    ctx.reset_accum(); ctx.render(0, 6); d = ctx.read_accum(False)
    ctx.set_options(samples_per_pass=0)
    assert np.allclose(a, b, rtol=1e-6, atol=1e-6) and np.allclose(a, d, rtol=1e-6, atol=1e-6)
    ctx.reset_accum(); ctx.render(0, 6); e = ctx.read_accum(False)
    assert a.tobytes() == e.tobytes()                                          # run-to-run deterministic


def test_stats_count_real_rays(ctx, get_case):
    c = get_case("cornell")
    load(ctx, c, api.TRAVERSAL_WIDE, max_depth=5, seed=1234, collect_stats=1)
    ctx.reset_accum(); ctx.reset_stats()
    ctx.render(0, 4)
    st = ctx.stats()
    _, _, orays = c.osc.render_pt(c.cam, 5, 1234, 0, 4)
    assert st["samples"] == c.w * c.h * 4
    # zero-weight paths are dropped instead of traced to the end, so fewer rays than the reference, never more
    assert 0.5 * orays < st["rays"] <= orays
    assert st["nodes_visited"] > st["rays"] and st["tris_tested"] > 0 and st["kernel_launches"] >= 12
    ctx.set_options(collect_stats=0)


def test_full_size_properties(ctx, lib):
    """config 2 at full size (262 144 triangles, 1024x1024): size-independent properties instead of the oracle"""
    sc = scenes.atrium()
    s = api.SceneHandle(); s.add_scene(sc)
    n2, d2 = s.build_bvh2()
    rank = api.bvh2_leaf_order(n2, s.counts()[1])
    b = s.build_bvh(tie_rank=rank)
    cam = api.make_camera(1024, 1024, **sc.camera)
    ctx.set_scene(s); ctx.set_bvh(b); ctx.set_bvh2(n2, d2); ctx.set_camera(cam)
    ctx.set_options(traversal=api.TRAVERSAL_WIDE, max_depth=5, seed=1234)
    wi, wt = ctx.primary_hits()
    ctx.set_options(traversal=api.TRAVERSAL_REF_BVH2)
    ri, rt = ctx.primary_hits()
    differ = (wi != ri) | (wt.view(np.uint32) != rt.view(np.uint32))
    assert differ.sum() <= 64 and np.all(wt[differ] <= rt[differ])            # closed room: wide never finds a farther hit
    # closed hall: apart from a handful of rays that slip between two triangles of the non-watertight test
    # (modelLoader.h:74-79; the brute force misses them too), every primary ray hits within the hall's diagonal
    assert (wi < 0).sum() <= 16 and np.all(wt[wi >= 0] < 60)
    assert (ri < 0).sum() <= 64                                                # (the reference's slab test may, BVH.h:65,79)
    # hit point lies on the reported triangle's plane
    rays = ctx.camera_rays()
    v, t, m, l = s.arrays()
    pos = v.view(np.float32).reshape(-1, 3).astype(np.float64)
    ok = wi >= 0
    p = rays[ok, :3] + rays[ok, 3:] * wt[ok, None].astype(np.float64)
    tr = t[wi[ok]]
    dist = np.einsum("ij,ij->i", p - pos[tr["v0"]], tr["norm"].astype(np.float64))
    assert np.abs(dist).max() < 1e-3
    # linearity of the accumulator: 2 x render(1 spp ranges) sums
    ctx.set_options(traversal=api.TRAVERSAL_WIDE)
    ctx.reset_accum(); ctx.render(0, 2); a = ctx.read_accum(False)
    ctx.reset_accum(); ctx.render(0, 1); ctx.render(1, 1); b2 = ctx.read_accum(False)
    assert np.allclose(a, b2, rtol=1e-6, atol=1e-6)
    # (the estimator is unbounded above: kernel.cu:504-506 multiplies by G = cos*cos/d^2 times the total light area)
    assert np.isfinite(a).all() and a.min() >= 0 and a.mean() > 0.1


def test_spheres_config1(ctx, get_case):
    """BASELINE config 1 (Cornell box + two spheres through sphere.h).  PARITY UNPINNED: the reference has the POD only
    (sphere.h:7-13, include commented out at kernel.cu:21); the checker is the oracle's restatement of the product's
    own sphere arithmetic.  Hits are compared bit for bit, radiance with the REF_BVH2-class tolerance (the Cornell box
    has no reference-dropped triangles)."""
    c = get_case("cornell_spheres")
    load(ctx, c, api.TRAVERSAL_WIDE, max_depth=5, seed=1234)
    ids, ts = ctx.primary_hits()
    oi, ot = c.osc.primary_hits(c.cam)
    assert np.array_equal(ids, oi) and ts.tobytes() == ot.tobytes()
    assert (ids < -1).sum() > 100                                           # both spheres are in view
    rng = np.random.default_rng(17)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    o = rng.uniform(pos.min(0), pos.max(0), size=(20000, 3)); d = rng.normal(size=(20000, 3))
    rays = np.concatenate([o, d / np.linalg.norm(d, axis=1, keepdims=True)], axis=1).astype(np.float32)
    assert hits_equal(ctx.trace_rays(rays), c.osc.brute_force(rays, c.rank))
    spp = 16                                                                # config 1: 16 spp, depth 5, seed 1234
    _, osum, _ = c.osc.render_pt(c.cam, 5, 1234, 0, spp)
    ref = osum / spp
    ctx.reset_accum(); ctx.render(0, spp)
    img = ctx.read_accum(True).astype(np.float64)
    close = np.all(np.abs(img - ref) <= 1e-3 * (1 + np.abs(ref)), axis=1)
    rel = np.sqrt(((img - ref) ** 2).mean()) / ref.mean()
    assert close.mean() >= 0.995 and rel <= 0.02, (close.mean(), rel)


@pytest.mark.parametrize("name", CASES)
def test_strict_mode_equals_the_reference_on_every_ray(ctx, get_case, name):
    """CUPT_TRAVERSAL_WIDE_STRICT: wide traversal + ancestor check in the reference tree + rare re-trace.  Hits are the
    reference's bit for bit on every ray (not only where the reference's slab test kept the nearest triangle), and
    the radiance image is bit-identical to the REF_BVH2 mode's, hence within 1e-6 of the oracle."""
    c = get_case(name)
    load(ctx, c, api.TRAVERSAL_WIDE_STRICT, max_depth=5, seed=1234)
    assert hits_equal(ctx.primary_hits(), c.osc.primary_hits(c.cam))
    rng = np.random.default_rng(23)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    o = rng.uniform(pos.min(0), pos.max(0), size=(30000, 3)); d = rng.normal(size=(30000, 3))
    rays = np.concatenate([o, d / np.linalg.norm(d, axis=1, keepdims=True)], axis=1).astype(np.float32)
    assert hits_equal(ctx.trace_rays(rays), c.osc.trace_rays(rays))
    ctx.reset_accum(); ctx.render(0, 4); strict = ctx.read_accum(False)
    ctx.set_options(traversal=api.TRAVERSAL_REF_BVH2)
    ctx.reset_accum(); ctx.render(0, 4); ref = ctx.read_accum(False)
    assert strict.tobytes() == ref.tobytes()


def test_strict_mode_full_size(ctx, lib):
    """config 2 at full size: the 26 primary rays on which the plain wide mode differs from the reference are gone"""
    sc = scenes.atrium()
    s = api.SceneHandle(); s.add_scene(sc)
    n2, d2 = s.build_bvh2()
    b = s.build_bvh(tie_rank=api.bvh2_leaf_order(n2, s.counts()[1]))
    ctx.set_scene(s); ctx.set_bvh(b); ctx.set_bvh2(n2, d2); ctx.set_camera(api.make_camera(1024, 1024, **sc.camera))
    ctx.set_options(traversal=api.TRAVERSAL_REF_BVH2, max_depth=5, seed=1234)
    ri, rt = ctx.primary_hits()
    ctx.set_options(traversal=api.TRAVERSAL_WIDE)
    wi, wt = ctx.primary_hits()
    ctx.set_options(traversal=api.TRAVERSAL_WIDE_STRICT)
    ctx.reset_stats()
    si, st = ctx.primary_hits()
    assert hits_equal((si, st), (ri, rt))
    plain_diff = int(((wi != ri) | (wt.view(np.uint32) != rt.view(np.uint32))).sum())
    assert 0 < plain_diff <= 64 and ctx.stats()["strict_retraces"] >= plain_diff
    ctx.reset_accum(); ctx.render(0, 2); a = ctx.read_accum(False)
    ctx.set_options(traversal=api.TRAVERSAL_REF_BVH2)
    ctx.reset_accum(); ctx.render(0, 2); r = ctx.read_accum(False)
    assert a.tobytes() == r.tobytes()
    ctx.set_options(traversal=api.TRAVERSAL_WIDE)
    # the live integrator (BDPT) at full size: strict == reference-exact mode bit for bit, the plain wide mode within noise
    ctx.set_options(integrator=api.INTEGRATOR_BDPT, traversal=api.TRAVERSAL_WIDE_STRICT, collect_stats=1)
    ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 1); bs = ctx.read_accum(False); rays_s = ctx.stats()["rays"]
    ctx.set_options(traversal=api.TRAVERSAL_REF_BVH2)
    ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 1); br = ctx.read_accum(False); rays_r = ctx.stats()["rays"]
    assert bs.tobytes() == br.tobytes() and rays_s == rays_r and 3.0 <= rays_r / (1024 * 1024) <= 7.0   # 3 sub-path rays + <= 4 connections
    ctx.set_options(traversal=api.TRAVERSAL_WIDE)
    ctx.reset_accum(); ctx.render(0, 1); bw = ctx.read_accum(False)
    assert (np.abs(bw - br).max(axis=1) > 1e-3 * (1 + np.abs(br).max(axis=1))).mean() < 1e-3 and abs(bw.mean() - br.mean()) < 0.01 * br.mean()
    ctx.set_options(integrator=api.INTEGRATOR_PT, collect_stats=0)
    # the device-built tree at full size: hits and image bit-identical to the host-built tree's; shared primary hits too
    ctx.reset_accum(); ctx.render(0, 2); host_img = ctx.read_accum(False)
    info = ctx.build_bvh_device(tie_rank=api.bvh2_leaf_order(n2, s.counts()[1]))
    assert info["num_tri_records"] == 262144 and info["depth"] <= 12
    di, dt = ctx.primary_hits()
    assert hits_equal((di, dt), (wi, wt))
    ctx.reset_accum(); ctx.render(0, 2)
    assert ctx.read_accum(False).tobytes() == host_img.tobytes()
    ctx.set_options(share_primary_hits=1)
    ctx.reset_accum(); ctx.render(0, 2)
    assert ctx.read_accum(False).tobytes() == host_img.tobytes()
    ctx.set_options(share_primary_hits=0)


def test_native_render_loop_example(ctx, lib, tmp_path):
    """examples/render_loop.cpp = the reference's main() (kernel.cu:565-790) in C++ over the C-ABI: OBJ in, PPM out.
    Its image must be byte-identical to the same render driven through the Python binding."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "examples", "render_loop")
    if not os.path.exists(exe):
        import __graft_entry__
        __graft_entry__.build()
    sc = scenes.cornell_box()
    obj, mtl = scenes.to_obj(sc)
    (tmp_path / "models").mkdir()
    (tmp_path / "models" / "cornell.obj").write_text(obj)
    (tmp_path / "models" / "scene.mtl").write_text(mtl)
    out = tmp_path / "image.ppm"
    r = subprocess.run([exe, str(tmp_path / "models" / "cornell.obj"), str(tmp_path / "models") + "/", "64", "48", "8", "5", str(out), "strict"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0, r.stdout
    assert "Mrays/s (counted)" in r.stdout and "32 tris" in r.stdout
    s = api.SceneHandle()
    s.load_obj(str(tmp_path / "models" / "cornell.obj"), str(tmp_path / "models") + "/")
    n2, d2 = s.build_bvh2()
    b = s.build_bvh(tie_rank=api.bvh2_leaf_order(n2, s.counts()[1]))
    ctx.set_scene(s); ctx.set_bvh(b); ctx.set_bvh2(n2, d2); ctx.set_camera(api.make_camera(64, 48, **sc.camera))
    ctx.set_options(traversal=api.TRAVERSAL_WIDE_STRICT, max_depth=5, seed=1234)
    ctx.reset_accum(); ctx.render(0, 8)
    ref = tmp_path / "ref.ppm"
    api.write_ppm(str(ref), ctx.read_accum(True), 64, 48)
    assert out.read_bytes() == ref.read_bytes()
    ctx.set_options(traversal=api.TRAVERSAL_WIDE)


def test_reference_style_main_over_the_compat_shim(ctx, lib, tmp_path):
    """examples/reference_main.cpp — the reference's main() shape (loadOBJ with its hard-coded "models/" MTL path, buildBVH,
    camera, depth guard, NUM_SAMPLES - 1 samples, the PPM written with the reference's own normalized / gammaCorrect) over
    include/cupt_compat.h — writes the same PPM, byte for byte, as examples/render_loop on the same scene."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe, exe2 = os.path.join(root, "examples", "reference_main"), os.path.join(root, "examples", "render_loop")
    if not os.path.exists(exe):
        pytest.skip("examples/reference_main is built where the reference's headers are (the container), and travels with the snapshot")
    sc = scenes.cornell_box()
    obj, mtl = scenes.to_obj(sc)
    (tmp_path / "models").mkdir()
    (tmp_path / "models" / "CornellBox-Original.obj").write_text(obj)
    (tmp_path / "models" / "scene.mtl").write_text(mtl)
    r = subprocess.run([exe, "models/CornellBox-Original.obj", "64", "48", "9", "shim.ppm"], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r.returncode == 0 and "32 tris" in r.stdout and "bvh size 31" in r.stdout, r.stdout
    r2 = subprocess.run([exe2, "models/CornellBox-Original.obj", "models/", "64", "48", "8", "3", "loop.ppm", "wide"], cwd=tmp_path, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=120)
    assert r2.returncode == 0, r2.stdout
    assert (tmp_path / "shim.ppm").read_bytes() == (tmp_path / "loop.ppm").read_bytes()


@pytest.mark.parametrize("name", CASES)
def test_bdpt_live_integrator_vs_oracle_and_golden(ctx, get_case, name):
    """CUPT_INTEGRATOR_BDPT = radianceAlongSingleStep (kernel.cu:217-415), what drawPixel calls at HEAD, as three
    extend rounds.  Same tolerances as the path tracer: the reference-exact traversal modes within 1e-3*(1+ref) on
    >= 99.9 % of the pixels and 1 % relative RMSE, the plain wide mode >= 98 % / 15 % / mean within 3 %.  The ray
    count is the reference's (3 sub-path rays + one per connection with non-zero weight), counted the same way."""
    c = get_case(name)
    g = np.load(os.path.join(GOLD, name + ".npz"))
    spp, seed = 4, 1234
    omean, osum, orays = c.osc.render_bdpt(c.cam, seed, 0, spp)
    assert osum.tobytes() == g["bdpt_sum_spp4_seed1234"].tobytes()
    want = osum / spp
    images = {}
    for mode, frac, rel_tol in ((api.TRAVERSAL_REF_BVH2, 0.999, 0.01), (api.TRAVERSAL_WIDE_STRICT, 0.999, 0.01), (api.TRAVERSAL_WIDE, 0.98, 0.15)):
        load(ctx, c, mode, integrator=api.INTEGRATOR_BDPT, seed=seed, collect_stats=1)
        ctx.reset_accum(); ctx.reset_stats()
        ctx.render(0, spp)
        img = ctx.read_accum(True).astype(np.float64)
        images[mode] = ctx.read_accum(False)
        st = ctx.stats()
        close = np.all(np.abs(img - want) <= 1e-3 * (1 + np.abs(want)), axis=1)
        rel = np.sqrt(((img - want) ** 2).mean()) / max(want.mean(), 1e-9)
        assert close.mean() >= frac, (name, mode, close.mean())
        assert rel <= rel_tol, (name, mode, rel)
        assert abs(img.mean() - want.mean()) <= 0.03 * want.mean(), (name, mode, img.mean(), want.mean())
        assert st["samples"] == c.w * c.h * spp
        assert abs(st["rays"] - orays) <= (0 if mode != api.TRAVERSAL_WIDE else 0.002 * orays), (name, mode, st["rays"], orays)
    assert images[api.TRAVERSAL_REF_BVH2].tobytes() == images[api.TRAVERSAL_WIDE_STRICT].tobytes()
    ctx.set_options(integrator=api.INTEGRATOR_PT, collect_stats=0)


def test_bdpt_sample_ranges_compose_and_reject_spheres(ctx, get_case):
    c = get_case("cornell_lights")
    load(ctx, c, api.TRAVERSAL_WIDE, integrator=api.INTEGRATOR_BDPT, seed=5)
    ctx.reset_accum(); ctx.render(0, 5); a = ctx.read_accum(False)
    ctx.set_options(samples_per_pass=2)
    ctx.reset_accum(); ctx.render(0, 3); ctx.render(3, 2); b = ctx.read_accum(False)
    ctx.set_options(samples_per_pass=0)
    assert np.allclose(a, b, rtol=1e-6, atol=1e-6)
    s = get_case("cornell_spheres")
    load(ctx, s, api.TRAVERSAL_WIDE, integrator=api.INTEGRATOR_BDPT)
    with pytest.raises(api.CuptError):
        ctx.render(0, 1)
    ctx.set_options(integrator=api.INTEGRATOR_PT)


def _random_rays(c, n, seed):
    rng = np.random.default_rng(seed)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    o = rng.uniform(pos.min(0), pos.max(0), size=(n, 3)); d = rng.normal(size=(n, 3))
    return np.concatenate([o, d / np.linalg.norm(d, axis=1, keepdims=True)], axis=1).astype(np.float32)


@pytest.mark.parametrize("name", CASES)
def test_device_bvh_build(ctx, get_case, name):
    """cupt_build_bvh_device (Morton sort + PLOC + 8-wide collapse on the GPU): the image it leaves in HBM satisfies the
    same structural invariants as the host builder's, builds are deterministic, and every hit — hence the radiance
    image — is bit-identical to the one traced through the host-built tree and to the oracle's brute force."""
    from bvhcheck import check_wide_bvh
    c = get_case(name)
    load(ctx, c, api.TRAVERSAL_WIDE, max_depth=5, seed=1234)
    host_hits = ctx.primary_hits()
    rays = _random_rays(c, 30000, 41)
    host_inco = ctx.trace_rays(rays)
    ctx.reset_accum(); ctx.render(0, 3); host_img = ctx.read_accum(False)
    for max_leaf in (1, 3):
        info = ctx.build_bvh_device(max_leaf_tris=max_leaf, tie_rank=c.rank)
        nodes, prims, depth, rounds = ctx.read_bvh()
        assert depth == info["depth"] and len(nodes) == info["num_nodes"] and rounds >= 1
        check_wide_bvh(c, nodes, prims, info, max_leaf)
        ctx.build_bvh_device(max_leaf_tris=max_leaf, tie_rank=c.rank)
        n2, p2, _, _ = ctx.read_bvh()
        assert nodes.tobytes() == n2.tobytes() and prims.tobytes() == p2.tobytes()      # deterministic
        assert hits_equal(ctx.primary_hits(), host_hits)
        assert hits_equal(ctx.trace_rays(rays), host_inco)
    bi, bt = c.osc.brute_force(rays, c.rank)
    assert hits_equal(host_inco, (bi, bt))
    ctx.reset_accum(); ctx.render(0, 3)
    assert ctx.read_accum(False).tobytes() == host_img.tobytes()
    # the strict mode works on top of the device-built tree as well (set_bvh2 was done by load())
    ctx.set_options(traversal=api.TRAVERSAL_WIDE_STRICT)
    assert hits_equal(ctx.primary_hits(), c.osc.primary_hits(c.cam))
    ctx.set_options(traversal=api.TRAVERSAL_WIDE)


def test_device_bvh_build_degenerate_inputs(ctx, get_case):
    """one triangle, two triangles, 3000 copies of the same triangle (all Morton codes and all boxes equal: the
    clustering must still halve every round, and the lowest rank wins every tie), spheres mixed with triangles"""
    c = get_case("cornell")
    pos = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 0, 1], [0, 1, 1]], np.float32)
    mats = c.mats[:1]
    for ntri in (1, 2, 3000):
        tris = np.zeros(ntri, dtype=c.tris.dtype)
        tris["v0"], tris["v1"], tris["v2"] = 0, 1, 2
        if ntri == 2:
            tris["v0"][1], tris["v1"][1], tris["v2"][1] = 3, 4, 5
        tris["norm"] = (0, 0, 1)
        s = api.SceneHandle()
        s.set_arrays(pos.view(c.verts.dtype).reshape(-1), tris, mats, np.zeros(0, np.int32), 0.0)
        ctx.set_scene(s)
        info = ctx.build_bvh_device()
        nodes, prims, depth, rounds = ctx.read_bvh()
        assert len(prims) == ntri and sorted(prims["id"].tolist()) == list(range(ntri))
        assert rounds <= 2 * int(np.ceil(np.log2(max(ntri, 2)))) + 2, rounds
        ctx.set_options(traversal=api.TRAVERSAL_WIDE)
        ray = np.array([[0.2, 0.2, 5, 0, 0, -1], [0.2, 0.2, -5, 0, 0, 1], [5, 5, 5, 0, 0, -1]], np.float32)
        ids, ts = ctx.trace_rays(ray)
        assert ids[2] == -1
        if ntri == 2:
            assert ids[0] == 1 and ts[0] == np.float32(4.0) and ids[1] == 0 and ts[1] == np.float32(5.0)
        else:
            assert ids[0] == 0 and ids[1] == 0 and ts[0] == np.float32(5.0)           # every copy ties: rank = index, lowest wins
    s = get_case("cornell_spheres")
    ctx.set_scene(s.handle); ctx.set_camera(s.cam)
    ctx.set_bvh(s.handle.build_bvh(tie_rank=s.rank))
    ctx.set_options(traversal=api.TRAVERSAL_WIDE)
    want = ctx.primary_hits()
    info = ctx.build_bvh_device(tie_rank=s.rank)
    assert info["num_spheres"] == len(s.spheres) and hits_equal(ctx.primary_hits(), want)


def test_device_output_stage(ctx, get_case, tmp_path):
    """tonemap on the device == the host restatement of kernel.cu:763-778 (which tests/
    test_scene_ingest.py pins against the reference's normalized/gammaCorrect) except where device pow differs from
    glibc's by an ulp right at a truncation boundary (<= 0.1 % of the values, by one level); heat map of the primary rays."""
    c = get_case("cornell_lights")
    load(ctx, c, api.TRAVERSAL_WIDE, max_depth=5, seed=3, collect_stats=1)
    ctx.reset_accum(); ctx.render(0, 8)
    want = api.tonemap(ctx.read_accum(True)).reshape(c.h, c.w, 3)
    got = ctx.read_image_rgb8()
    diff = np.abs(got.astype(np.int32) - want.astype(np.int32))
    assert got.shape == (c.h, c.w, 3) and diff.max() <= 1 and (diff != 0).mean() <= 1e-3
    assert np.array_equal(ctx.read_image_rgb8(mirror_x=True), got[:, ::-1])
    api.write_ppm8(str(tmp_path / "a.ppm"), got[:, ::-1]); api.write_ppm(str(tmp_path / "b.ppm"), ctx.read_accum(True), c.w, c.h)
    ta, tb = open(tmp_path / "a.ppm").read().split(), open(tmp_path / "b.ppm").read().split()
    assert ta[:4] == tb[:4] and len(ta) == len(tb) and np.mean(np.array(ta[4:], int) != np.array(tb[4:], int)) <= 1e-3
    nodes, prims = ctx.read_heatmap()
    ids, _ = ctx.primary_hits()
    assert nodes.shape == (c.h, c.w) and nodes.min() >= 1 and prims[ids.reshape(c.h, c.w) >= 0].min() >= 1
    ctx.reset_stats(); ctx.primary_hits(); st = ctx.stats()
    # the wavefront kernel defers primitive tests by one step, so it culls a little later than the single-ray walk
    assert 0.9 * st["nodes_visited"] <= nodes.sum() <= 1.1 * st["nodes_visited"]
    assert 0.9 * st["tris_tested"] <= prims.sum() <= 1.1 * st["tris_tested"]
    ctx.set_options(collect_stats=0)


def test_watertight_triangle_test(ctx, get_case):
    """CUPT_TRI_WATERTIGHT (north_star: "watertight ray-triangle tests"): rays aimed at shared edges and vertices of a
    tessellated, rotated plane never fall through, whereas the reference's own test (modelLoader.h:49-82, the parity
    mode) lets a few slip; away from edges the two tests agree (same triangle, t within 1e-5 relative).  Outside the
    bit-exact parity statement by construction: the tolerance here is the watertight property itself."""
    c = get_case("cornell")
    rng = np.random.default_rng(11)
    n = 40
    gx = np.cumsum(rng.uniform(0.5, 1.5, n + 1)); gy = np.cumsum(rng.uniform(0.5, 1.5, n + 1))
    gx = (gx - gx.mean()) / 8; gy = (gy - gy.mean()) / 8
    P = np.stack(np.meshgrid(gx, gy, indexing="ij"), -1).reshape(-1, 2)
    q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
    pos = (np.concatenate([P, np.zeros((len(P), 1))], 1) @ q.T).astype(np.float32)
    idx = lambda i, j: i * (n + 1) + j
    tri = []
    for i in range(n):
        for j in range(n):
            tri += [(idx(i, j), idx(i + 1, j), idx(i + 1, j + 1)), (idx(i, j), idx(i + 1, j + 1), idx(i, j + 1))]
    tris = np.zeros(len(tri), dtype=c.tris.dtype)
    tris["v0"], tris["v1"], tris["v2"] = np.array(tri).T
    tris["norm"] = q[:, 2].astype(np.float32)
    s = api.SceneHandle()
    s.set_arrays(pos.view(c.verts.dtype).reshape(-1), tris, c.mats[:1], np.zeros(0, np.int32), 0.0)
    ctx.set_scene(s); ctx.build_bvh_device()
    # targets on interior edges (including the diagonals) and on interior vertices, in f32 like the mesh
    ii, jj = np.meshgrid(np.arange(1, n - 1), np.arange(1, n - 1), indexing="ij")
    a = pos[idx(ii, jj).ravel()]
    targets = [a]
    for di, dj in ((1, 0), (0, 1), (1, 1)):
        b = pos[idx(ii + di, jj + dj).ravel()]
        for w in (0.25, 0.5, 0.8125):
            targets.append((a + np.float32(w) * (b - a)).astype(np.float32))
    targets = np.concatenate(targets)
    origin = (q[:, 2] * 7.0 + q[:, 0] * 1.3).astype(np.float32)
    d = targets - origin
    d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    rays = np.concatenate([np.broadcast_to(origin, d.shape), d], 1).astype(np.float32)
    ctx.set_options(traversal=api.TRAVERSAL_WIDE, tri_test=api.TRI_REFEXACT)
    ri, rt = ctx.trace_rays(rays)
    ctx.set_options(tri_test=api.TRI_WATERTIGHT)
    wi, wt = ctx.trace_rays(rays)
    assert (wi >= 0).all(), int((wi < 0).sum())                                 # nothing falls through
    slipped = int((ri < 0).sum())
    both = (ri >= 0)
    assert np.all(np.abs(wt[both] - rt[both]) <= 1e-5 * rt[both])
    # interior targets: same triangle
    ctr = ((pos[tris["v0"]] + pos[tris["v1"]] + pos[tris["v2"]]) / np.float32(3)).astype(np.float32)
    d = ctr - origin; d = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
    rays2 = np.concatenate([np.broadcast_to(origin, d.shape), d], 1).astype(np.float32)
    wi2, wt2 = ctx.trace_rays(rays2)
    assert np.array_equal(wi2, np.arange(len(tris)))
    # strict / reference modes refuse the non-reference test; a real scene renders to the same image within noise
    ctx.set_options(traversal=api.TRAVERSAL_REF_BVH2)
    with pytest.raises(api.CuptError):
        ctx.trace_rays(rays2)
    ctx.set_options(traversal=api.TRAVERSAL_WIDE, tri_test=api.TRI_REFEXACT)
    a20 = get_case("atrium20k")
    load(ctx, a20, api.TRAVERSAL_WIDE, max_depth=5, seed=9)
    ri, rt = ctx.primary_hits(); ctx.reset_accum(); ctx.render(0, 4); ref_img = ctx.read_accum(True)
    ctx.set_options(tri_test=api.TRI_WATERTIGHT)
    wi, wt = ctx.primary_hits(); ctx.reset_accum(); ctx.render(0, 4); wt_img = ctx.read_accum(True)
    ctx.set_options(tri_test=api.TRI_REFEXACT)
    # coincident faces (boxes standing on the floor) tie in t and the winner then depends on the last ulp: the triangle
    # may differ there, the surface point may not
    both = (ri >= 0) & (wi >= 0)
    assert (ri == wi).mean() >= 0.99 and (wi < 0).sum() <= (ri < 0).sum()
    assert np.all(np.abs(wt[both] - rt[both]) <= 2e-5 * rt[both])
    assert abs(wt_img.mean() - ref_img.mean()) <= 0.02 * ref_img.mean()
    print("rays through shared edges lost by the reference test:", slipped, "of", len(rays))


def test_specular_materials_config5(ctx, get_case):
    """BASELINE config 5: Cornell box with triangle lights and mixed diffuse / mirror materials at max depth 16.  The
    mirror is an extension (the reference has diffuse materials only): PARITY UNPINNED, checked against the oracle's
    restatement of the same rule (d - 2 (d.n) n, weight *= albedo, no random numbers) with the path tracer's tolerances."""
    c = get_case("cornell_specular")
    plain = get_case("cornell_lights")
    spp, seed, nb = 4, 21, 16
    _, osum, orays = c.osc.render_pt(c.cam, nb, seed, 0, spp)
    _, psum, _ = plain.osc.render_pt(plain.cam, nb, seed, 0, spp)
    want = osum / spp
    assert np.abs(osum - psum).max() > 0.1                                            # the mirrors change the image
    for mode, frac, rel_tol in ((api.TRAVERSAL_REF_BVH2, 0.999, 0.01), (api.TRAVERSAL_WIDE, 0.98, 0.15)):
        load(ctx, c, mode, max_depth=nb, seed=seed, collect_stats=1)
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, spp)
        img = ctx.read_accum(True).astype(np.float64)
        close = np.all(np.abs(img - want) <= 1e-3 * (1 + np.abs(want)), axis=1)
        rel = np.sqrt(((img - want) ** 2).mean()) / want.mean()
        assert close.mean() >= frac and rel <= rel_tol, (mode, close.mean(), rel)
        assert 0.3 * orays < ctx.stats()["rays"] <= orays
    ctx.set_options(collect_stats=0, max_depth=5)


@pytest.mark.parametrize("name", ["cornell_lights", "atrium20k"])
def test_shared_primary_hits(ctx, get_case, name):
    """share_primary_hits: with the reference's pinhole camera (radius 0, kernel.cu:646) the primary ray is the same
    for every sample of a pixel; tracing it once per pixel and pass gives the same image bit for bit with fewer rays.
    With a lens (radius > 0) the option must switch itself off."""
    c = get_case(name)
    for mode in (api.TRAVERSAL_WIDE, api.TRAVERSAL_WIDE_STRICT):
        # (the reference-exact modes do not share: on the centre column the direction's x is +0 or -0 depending on the
        # sample's lens angle, and the reference's slab test, which divides by it, returns different hits for the two)
        saved = 1 if mode == api.TRAVERSAL_WIDE else 0
        load(ctx, c, mode, max_depth=5, seed=77, collect_stats=1, share_primary_hits=0)
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 6); a = ctx.read_accum(False); ra = ctx.stats()["rays"]
        ctx.set_options(share_primary_hits=1)
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 6); b = ctx.read_accum(False); rb = ctx.stats()["rays"]
        assert a.tobytes() == b.tobytes()
        assert rb == ra - saved * 5 * c.w * c.h                                    # five of six primary rays per pixel saved
        ctx.set_options(samples_per_pass=4)                                        # passes of 4 + 2 samples
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 6); d = ctx.read_accum(False)
        assert np.allclose(a, d, rtol=1e-6, atol=1e-6) and ctx.stats()["rays"] == ra - saved * 4 * c.w * c.h
        ctx.set_options(samples_per_pass=0)
    lens = api.make_camera(c.w, c.h, **dict(c.scene.camera, radius=0.05))
    ctx.set_camera(lens)
    ctx.set_options(traversal=api.TRAVERSAL_WIDE, share_primary_hits=0)
    ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 3); a = ctx.read_accum(False); ra = ctx.stats()["rays"]
    ctx.set_options(share_primary_hits=1)
    ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 3); b = ctx.read_accum(False)
    assert a.tobytes() == b.tobytes() and ctx.stats()["rays"] == ra
    ctx.set_options(share_primary_hits=0, collect_stats=0)


def test_device_radix_sort(ctx):
    """the hand-written stable LSD radix sort behind the device BVH build (csrc/radix_sort.cuh) against numpy's stable
    argsort: ragged sizes around the 4096-pair tile, heavy duplicates, partial-key sorts (top 8 * passes bits)"""
    rng = np.random.default_rng(3)
    for n in (1, 31, 4095, 4096, 4097, 100003, 1 << 20):
        keys = rng.integers(0, 1 << 63, n, dtype=np.uint64)
        keys[::3] = keys[0]                                                      # a third of the keys are equal
        keys[1::7] &= np.uint64(0xFFFFFF0000000000)                              # and some differ only in the top bits
        vals = np.arange(n, dtype=np.uint32)
        for passes in (8, 5, 1):
            shift = max(0, 63 - 8 * passes)
            want = np.argsort(keys >> np.uint64(shift), kind="stable")
            k, v = ctx.debug_sort_pairs(keys, vals, passes)
            assert np.array_equal(v, want.astype(np.uint32)), (n, passes)
            assert np.array_equal(k, keys[want])


def test_large_scene_upload_and_soup_hits(ctx, lib):
    """a 1.5 M-triangle soup (config 4's generator): its arrays (42 MB of triangles, 72 MB of primitive records) outgrow the
    32 MB pinned staging buffer of the uploads (context.cu: stage_alloc), so the grow / rewind / wait paths run; hits
    through the host-built and the device-built tree against the oracle's brute force over ALL triangles, bit for bit,
    then a small scene again on the same context (the buffer is reused)"""
    import oraclebind
    n = 1_500_000
    sc = scenes.triangle_soup(n)
    s = api.SceneHandle(); s.add_scene(sc)
    v, t, mm, l = s.arrays()
    rng = np.random.default_rng(5)
    o = rng.uniform(-50, 50, size=(300, 3)); d = rng.normal(size=(300, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    osc = oraclebind.OracleScene(v, t, mm, l, s.total_light_area, np.zeros(1, api.BVH2_NODE))
    want = osc.brute_force(rays, None)
    assert (want[0] >= 0).sum() > 30
    ctx.set_options(traversal=api.TRAVERSAL_WIDE)
    for _ in range(2):                                                           # twice: first use grows the buffer, second reuses it
        ctx.set_scene(s)
        b = s.build_bvh()
        ctx.set_bvh(b)
        assert hits_equal(ctx.trace_rays(rays), want)
        b.close()
    ctx.set_scene(s)
    ctx.build_bvh_device()
    assert hits_equal(ctx.trace_rays(rays), want)
    s.close()
    # and a 12-triangle scene right after it
    small = scenes.cornell_box() if hasattr(scenes, "cornell_box") else None
    if small is not None:
        s2 = api.SceneHandle(); s2.add_scene(small)
        v2, t2, m2, l2 = s2.arrays()
        o2 = oraclebind.OracleScene(v2, t2, m2, l2, s2.total_light_area, np.zeros(1, api.BVH2_NODE))
        r2 = np.concatenate([rng.uniform(-0.9, 0.9, size=(200, 3)) + [0, 1, 0], rng.normal(size=(200, 3))], axis=1).astype(np.float32)
        ctx.set_scene(s2)
        b2 = s2.build_bvh(); ctx.set_bvh(b2)
        assert hits_equal(ctx.trace_rays(r2), o2.brute_force(r2, None))
        b2.close(); s2.close()


def test_tile_windows_compose(ctx, get_case):
    """SURVEY §8e, the tile alternative to sample ranges: rendering the image as disjoint tile windows (8x4-pixel tiles,
    split like sharding.tile_range does for 3 ranks) gives the whole-image render bit for bit — every pixel gets the
    same Philox streams whichever window it is rendered in — and a window leaves the pixels outside it untouched;
    both integrators"""
    from cudapathtracer_b200 import sharding
    c = get_case("cornell")
    for integ in (api.INTEGRATOR_PT, api.INTEGRATOR_BDPT):
        load(ctx, c, api.TRAVERSAL_WIDE, max_depth=5, seed=1234, integrator=integ)
        ctx.set_tile_window(0, 0)
        ctx.reset_accum(); ctx.render(0, 3)
        whole = ctx.read_accum(False).copy()
        n = ctx.tile_count()
        w, h = int(c.cam["pxlWidth"][0]), int(c.cam["pxlHeight"][0])
        assert n == ((w + 7) // 8) * ((h + 3) // 4)
        ctx.reset_accum()
        parts = []
        for r in range(3):
            first, cnt = sharding.tile_range(r, 3, n)
            ctx.set_tile_window(first, cnt)
            ctx.render(0, 3)
            parts.append(ctx.read_accum(False).copy())
        assert parts[-1].tobytes() == whole.tobytes()                             # the windows accumulate into one image
        first, cnt = sharding.tile_range(0, 3, n)
        touched = (parts[0].reshape(h, w, 3) != 0).any(axis=2)
        ty, tx = np.divmod(np.arange(first, first + cnt), (w + 7) // 8)
        inside = np.zeros((h, w), bool)
        for y0, x0 in zip(ty * 4, tx * 8):
            inside[y0:y0 + 4, x0:x0 + 8] = True
        assert not touched[~inside].any() and touched[inside].mean() > 0.5        # nothing outside window 0 after the first part
        ctx.set_tile_window(n + 5, 4)
        with pytest.raises(Exception):
            ctx.render(0, 1)
        ctx.set_tile_window(0, 0)
    ctx.set_options(integrator=api.INTEGRATOR_PT)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cornell_lights", "atrium20k", "soup5k", "cornell_spheres", "cornell_specular"])
def test_last_segment_early_stop_changes_nothing(ctx, get_case, name):
    """The ray of a path's last segment may stop at the first hit in front of the emitters' bounding box (k_shade:
    emitter_entry_distance, k_extend's visibility variant): the next k_shade only asks that hit for its emission
    (kernel.cu:453-457), so the image must not move.  cupt_render_opts.no_terminal_stop = 1 turns the early stop off; the few pixels
    allowed to differ are rays grazing an emitter, where triIntersect's own rounding decides which side of the
    shortened distance its t falls on."""
    c = get_case(name)
    load(ctx, c, api.TRAVERSAL_WIDE, max_depth=5, seed=11, samples_per_pass=0, integrator=api.INTEGRATOR_PT, collect_stats=0)
    try:
        ctx.set_options(no_terminal_stop=1)
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 8); a = ctx.read_accum(False)
        ra = ctx.stats()["rays"]
        ctx.set_options(no_terminal_stop=0)
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 8); b = ctx.read_accum(False)
        rb = ctx.stats()["rays"]
    finally:
        ctx.set_options(no_terminal_stop=0)
    same = np.all(a == b, axis=-1)
    assert same.mean() >= 0.999, (name, same.mean())
    assert ra == rb > 0   # every ray is still traced, only some of them less far
    assert a.sum() > 0 and abs(float(a.sum()) - float(b.sum())) <= 1e-3 * float(a.sum())



@pytest.mark.parametrize("name", ["atrium20k", "soup5k"])
def test_extend_tuning_changes_no_hit_and_no_image(ctx, get_case, name):
    """fetch_threshold (when idle lanes are refilled) and defer_pairs (how many (ray, primitive) pairs a warp collects before a
    shared test round; lanes that cannot go on without their tests wait for it) schedule the work of k_extend's warps and must
    not show in any result: the closest hit is an argmin with a total tie-break order, whichever lane tests which pair when.
    1 pair = a round in every iteration, 64 = rounds only when nothing else can move (every lane waiting)."""
    c = get_case(name)
    rng = np.random.default_rng(17)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    o = rng.uniform(pos.min(0), pos.max(0), size=(40000, 3))
    d = rng.normal(size=(40000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    base = None
    try:
        for ft, dp in [(0, 0), (1, 1), (8, 1), (31, 1), (24, 8), (24, 31), (31, 64), (4, 64), (16, 33)]:
            for integ in (api.INTEGRATOR_PT, api.INTEGRATOR_BDPT):
                load(ctx, c, api.TRAVERSAL_WIDE, max_depth=5, seed=3, integrator=integ, fetch_threshold=ft, defer_pairs=dp, collect_stats=0)
                ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 4)
                got = (ctx.trace_rays(rays), ctx.read_accum(False).tobytes(), ctx.stats()["rays"], integ)
                if base is None or base[integ] is None:
                    base = base or {api.INTEGRATOR_PT: None, api.INTEGRATOR_BDPT: None}
                    base[integ] = got
                    continue
                ref = base[integ]
                assert hits_equal(got[0], ref[0]), (ft, dp, integ)
                assert got[1] == ref[1] and got[2] == ref[2], (ft, dp, integ)
    finally:
        ctx.set_options(fetch_threshold=0, defer_pairs=0, integrator=api.INTEGRATOR_PT)
    bi, bt = c.osc.brute_force(rays, c.rank)
    assert np.array_equal(base[api.INTEGRATOR_PT][0][0], bi) and base[api.INTEGRATOR_PT][0][1].tobytes() == bt.tobytes()
