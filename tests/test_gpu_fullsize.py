"""GPU parity at the sizes BASELINE.json names, against the ORACLE (oracle/oracle.c, pinned bit for bit to the compiled
reference by tests/test_oracle.py) — not GPU mode against GPU mode:

  config 2   atrium, 262 144 triangles, 1024 x 1024: primary hit ids and t bits of the reference-exact modes equal the
             oracle's trace(); the wide mode equals the oracle's brute force over ALL triangles, and every pixel where it
             differs from the reference is one the strict mode lists as "the reference would not have reached this hit";
             radiance of the three modes against the oracle's integrator, the bench's own mode (wide, last-segment early
             stop on) included
             the reference's LIVE integrator (BDPT, kernel.cu:217-415) at the same size: three modes against the oracle's
             restatement of it
  config 3   the same scene at 3840 x 2160: the only check of the 8x4 tile-major slot indexing on a non-square,
             non-power-of-two image (the reference's 16-bit Morton map, camera.h:57-65, cannot express it)
  config 1   Cornell box at 256 x 256, 16 spp, depth 5, seed 1234 — triangle-only (parity pinned) and with the two
             spheres (parity unpinned: sphere.h is a POD)

Tolerances.  Hits: bit-exact.  Radiance (f32 sums vs the reference's f64): paths are seed-matched, so apart from device
sinf/cosf ulps the images agree to rounding: >= 99.9 % of the pixels within 1e-3 * (1 + ref) and relative RMSE <= 1 % —
for the wide mode on the pixels none of whose rays the strict mode had to re-trace (there the reference itself lost
its nearest triangle, BVH.h:65,79, and the path continues differently).
"""
import numpy as np
import pytest

from cudapathtracer_b200 import api, scenes

pytestmark = pytest.mark.gpu


def hits_equal(a, b):
    return np.array_equal(a[0], b[0]) and a[1].tobytes() == b[1].tobytes()


def image_stats(img, want, keep=None):
    """(fraction of pixels within 1e-3 (1 + ref), relative RMSE, relative mean difference) over the pixels in `keep`"""
    img = np.asarray(img, np.float64); want = np.asarray(want, np.float64)
    if keep is not None:
        img, want = img[keep], want[keep]
    close = np.all(np.abs(img - want) <= 1e-3 * (1 + np.abs(want)), axis=1)
    rel = np.sqrt(((img - want) ** 2).mean()) / max(want.mean(), 1e-12)
    return close.mean(), rel, abs(img.mean() - want.mean()) / max(want.mean(), 1e-12)


class Full:
    """config 2's scene with everything the three traversal modes need, and its oracle twin"""

    def __init__(self, oraclebind):
        self.sc = scenes.atrium()
        self.s = api.SceneHandle(); self.s.add_scene(self.sc)
        self.v, self.t, self.m, self.l = self.s.arrays()
        self.n2, self.d2 = self.s.build_bvh2()
        self.rank = api.bvh2_leaf_order(self.n2, len(self.t))
        self.bvh = self.s.build_bvh(tie_rank=self.rank)
        self.osc = oraclebind.OracleScene(self.v, self.t, self.m, self.l, self.s.total_light_area, self.n2)

    def load(self, ctx, cam, mode, **opts):
        ctx.set_scene(self.s); ctx.set_bvh(self.bvh); ctx.set_bvh2(self.n2, self.d2); ctx.set_camera(cam)
        o = dict(traversal=mode, integrator=api.INTEGRATOR_PT, tri_test=api.TRI_REFEXACT, samples_per_pass=0,
                 share_primary_hits=0, no_terminal_stop=0, record_retraces=0, collect_stats=0)
        o.update(opts)
        ctx.set_options(**o)


@pytest.fixture(scope="module")
def full(lib, oracle):
    return Full(oracle)


@pytest.fixture(scope="module")
def ctx(lib):
    c = api.Context(0)
    yield c
    c.close()


def check_primary(ctx, full, cam, n_brute):
    """the three modes' primary hits against the oracle; returns the number of rays on which wide != reference"""
    oi, ot = full.osc.primary_hits(cam)
    full.load(ctx, cam, api.TRAVERSAL_REF_BVH2)
    assert hits_equal(ctx.primary_hits(), (oi, ot))
    full.load(ctx, cam, api.TRAVERSAL_WIDE_STRICT, record_retraces=1)
    ctx.reset_accum()
    assert hits_equal(ctx.primary_hits(), (oi, ot))
    listed = ctx.read_retrace_counts() > 0
    full.load(ctx, cam, api.TRAVERSAL_WIDE)
    wi, wt = ctx.primary_hits()
    differ = (wi != oi) | (wt.view(np.uint32) != ot.view(np.uint32))
    # (how many depends on how the pixel grid falls on the scene's box faces: 26 of 2^20 at 1024 x 1024, 2674 of 8.3 M at 4K)
    assert 0 < differ.sum() <= 1e-3 * len(oi)
    w = int(cam["pxlWidth"][0])
    ys, xs = np.divmod(np.flatnonzero(differ), w)
    print("rows with most differing rays:", sorted(zip(*np.unique(ys, return_counts=True)), key=lambda a: -a[1])[:4],
          "columns:", sorted(zip(*np.unique(xs, return_counts=True)), key=lambda a: -a[1])[:4])
    assert np.all(listed[differ]), "a ray on which wide != reference is missing from the strict mode's re-trace list"
    # the wide mode's answer is the argmin over ALL triangles: brute force on every differing ray and on a random sample
    rays = full.osc.camera_rays(cam)
    rng = np.random.default_rng(8)
    pick = np.unique(np.concatenate([np.flatnonzero(differ), rng.integers(0, len(oi), n_brute)]))
    bi, bt = full.osc.brute_force(rays[pick], full.rank)
    assert np.array_equal(wi[pick], bi) and wt[pick].tobytes() == bt.tobytes()
    assert np.all(wt[differ] <= ot[differ])
    return int(differ.sum())


def test_config2_primary_hits_vs_oracle(ctx, full):
    cam = api.make_camera(1024, 1024, **full.sc.camera)
    full.load(ctx, cam, api.TRAVERSAL_WIDE)
    assert ctx.camera_rays().tobytes() == full.osc.camera_rays(cam).tobytes()
    n = check_primary(ctx, full, cam, 1500)
    print("config 2: wide != reference on", n, "of", 1024 * 1024, "primary rays, all explained by the brute force")


def test_config3_primary_hits_vs_oracle(ctx, full):
    cam = api.make_camera(3840, 2160, **full.sc.camera)
    full.load(ctx, cam, api.TRAVERSAL_WIDE)
    assert ctx.camera_rays().tobytes() == full.osc.camera_rays(cam).tobytes()
    n = check_primary(ctx, full, cam, 1000)
    print("config 3: wide != reference on", n, "of", 3840 * 2160, "primary rays")


@pytest.mark.parametrize("spp,seed", [(1, 1234), (2, 77)])
def test_config2_radiance_vs_oracle(ctx, full, spp, seed):
    cam = api.make_camera(1024, 1024, **full.sc.camera)
    _, osum, orays = full.osc.render_pt(cam, 5, seed, 0, spp)
    want = osum / spp
    full.load(ctx, cam, api.TRAVERSAL_REF_BVH2, max_depth=5, seed=seed)
    ctx.reset_accum(); ctx.render(0, spp)
    f, rel, dm = image_stats(ctx.read_accum(True), want)
    print(f"config 2 {spp} spp REF_BVH2: {f:.6f} of the pixels within tolerance, rel. RMSE {rel:.2e}")
    assert f >= 0.999 and rel <= 0.01 and dm <= 0.001, (f, rel, dm)
    full.load(ctx, cam, api.TRAVERSAL_WIDE_STRICT, max_depth=5, seed=seed, record_retraces=1)
    ctx.reset_accum(); ctx.render(0, spp)
    strict = ctx.read_accum(True)
    f, rel, dm = image_stats(strict, want)
    assert f >= 0.999 and rel <= 0.01 and dm <= 0.001, (f, rel, dm)
    clean = ctx.read_retrace_counts() == 0
    assert clean.mean() >= 0.995                     # about one ray in 5 400 is re-traced
    # the bench's own mode: wide, last-segment early stop ON
    full.load(ctx, cam, api.TRAVERSAL_WIDE, max_depth=5, seed=seed, collect_stats=1)
    ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, spp)
    wide = ctx.read_accum(True)
    f, rel, dm = image_stats(wide, want, clean)
    print(f"config 2 {spp} spp WIDE (early stop on), pixels without a re-traced ray ({clean.mean():.4f} of all): {f:.6f} within tolerance, rel. RMSE {rel:.2e}")
    assert f >= 0.999 and rel <= 0.01 and dm <= 0.001, (f, rel, dm)
    f_all, rel_all, dm_all = image_stats(wide, want)
    assert f_all >= 0.995 and dm_all <= 0.01, (f_all, rel_all, dm_all)
    # rays: zero-weight paths are dropped instead of traced to the end, never more than the reference traces
    assert 0.5 * orays < ctx.stats()["rays"] <= orays
    ctx.set_options(collect_stats=0)


def test_config2_bdpt_vs_oracle(ctx, full):
    """The reference's LIVE integrator (radianceAlongSingleStep, kernel.cu:217-415 — what drawPixel calls at HEAD) at config 2's
    size: 1024 x 1024, 1 spp, seed 1234, against the oracle's restatement of it (bit-exact against the compiled reference,
    tests/test_oracle.py).  Reference-exact traversal modes: 99.9 % of the pixels within 1e-3 (1 + ref), relative RMSE 1 %, and the
    reference's ray count to 1e-5; the wide mode the same on the pixels none of whose rays the strict mode re-traced, its ray
    count within 0.2 % (a path that continues differently makes different connections)."""
    cam = api.make_camera(1024, 1024, **full.sc.camera)
    _, osum, orays = full.osc.render_bdpt(cam, 1234, 0, 1)
    want = osum
    imgs = {}
    for mode in (api.TRAVERSAL_REF_BVH2, api.TRAVERSAL_WIDE_STRICT):
        full.load(ctx, cam, mode, integrator=api.INTEGRATOR_BDPT, seed=1234, collect_stats=1, record_retraces=int(mode == api.TRAVERSAL_WIDE_STRICT))
        ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 1)
        imgs[mode] = ctx.read_accum(False)
        f, rel, dm = image_stats(imgs[mode], want)
        print(f"config 2 BDPT mode {mode}: {f:.6f} of the pixels within tolerance, rel. RMSE {rel:.2e}, rays {ctx.stats()['rays']} (reference {orays})")
        assert f >= 0.999 and rel <= 0.01 and dm <= 0.001, (mode, f, rel, dm)
        # (a connection gets its visibility ray when its weight exceeds 0.01, kernel.cu:396: a device sinf / cosf ulp moves a
        # handful of weights across that threshold — 2 of 5.8 M rays here; the small cases of test_gpu_parity.py agree exactly)
        assert abs(ctx.stats()["rays"] - orays) <= 1e-5 * orays
    assert imgs[api.TRAVERSAL_REF_BVH2].tobytes() == imgs[api.TRAVERSAL_WIDE_STRICT].tobytes()
    clean = ctx.read_retrace_counts() == 0
    assert clean.mean() >= 0.99
    full.load(ctx, cam, api.TRAVERSAL_WIDE, integrator=api.INTEGRATOR_BDPT, seed=1234, collect_stats=1)
    ctx.reset_accum(); ctx.reset_stats(); ctx.render(0, 1)
    wide = ctx.read_accum(False)
    f, rel, dm = image_stats(wide, want, clean)
    print(f"config 2 BDPT WIDE, pixels without a re-traced ray ({clean.mean():.4f} of all): {f:.6f} within tolerance, rel. RMSE {rel:.2e}")
    assert f >= 0.999 and rel <= 0.01 and dm <= 0.001, (f, rel, dm)
    f_all, _, dm_all = image_stats(wide, want)
    assert f_all >= 0.99 and dm_all <= 0.01, (f_all, dm_all)
    assert abs(ctx.stats()["rays"] - orays) <= 0.002 * orays
    ctx.set_options(integrator=api.INTEGRATOR_PT, collect_stats=0, record_retraces=0)


def test_config1_full_size(ctx, get_case, oracle):
    """BASELINE config 1: 256 x 256, 16 spp, max depth 5, seed 1234"""
    for name in ("cornell", "cornell_spheres"):
        c = get_case(name)
        cam = api.make_camera(256, 256, **c.scene.camera)
        oi, ot = c.osc.primary_hits(cam)
        _, osum, orays = c.osc.render_pt(cam, 5, 1234, 0, 16)
        want = osum / 16
        modes = (api.TRAVERSAL_REF_BVH2, api.TRAVERSAL_WIDE_STRICT, api.TRAVERSAL_WIDE) if name == "cornell" else (api.TRAVERSAL_WIDE,)
        for mode in modes:
            ctx.set_scene(c.handle)
            ctx.set_bvh(c.handle.build_bvh(tie_rank=c.rank)); ctx.set_bvh2(c.nodes2, c.depth2); ctx.set_camera(cam)
            ctx.set_options(traversal=mode, max_depth=5, seed=1234, integrator=api.INTEGRATOR_PT, record_retraces=0, collect_stats=0)
            assert hits_equal(ctx.primary_hits(), (oi, ot)), (name, mode)   # (the Cornell box has no reference-dropped triangles)
            ctx.reset_accum(); ctx.render(0, 16)
            f, rel, dm = image_stats(ctx.read_accum(True), want)
            print(f"config 1 {name} mode {mode}: {f:.6f} within tolerance, rel. RMSE {rel:.2e}")
            assert f >= 0.999 and rel <= 0.01 and dm <= 0.002, (name, mode, f, rel, dm)
