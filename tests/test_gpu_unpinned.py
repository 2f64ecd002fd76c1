"""Independent checks for the two rows the reference cannot pin (PARITY UNPINNED):

  * ray-sphere: sphere.h:7-13 is a POD, the reference has no intersector.  The oracle's sphere code restates the product's own
    operation order, so a wrong formula would pass both.  Here the checker is the textbook quadratic in float64 — a different
    formula, not the product's sequence (closest approach + the numerically stable root pair) — on the float32 rays the GPU traced:
    the hit distance must agree within 4 ulp where the problem is well-conditioned (4 ulp x its conditioning elsewhere) and the hit / miss decision must agree except within 1e-6 (relative to the radius)
    of tangency.
  * mirror materials (config 5's extension): with a mirror floor under a rectangular emitter and max depth 2 the image has a
    closed form — albedo x emission where the REFLECTED ray d - 2 (d.n) n meets the emitter, emission where the primary ray
    does, 0 elsewhere — evaluated here in float64.
"""
import numpy as np
import pytest

from cudapathtracer_b200 import api

pytestmark = pytest.mark.gpu
ULP = np.float64(2.0 ** -23)


@pytest.fixture(scope="module")
def ctx(lib):
    c = api.Context(0)
    yield c
    c.close()


def sphere_roots_f64(o, d, centre, radius):
    """nearest root t > 0 of |o + t d - c|^2 = r^2 per ray (inf if none) and the tangency margin (disc / (a r^2))"""
    o, d = o.astype(np.float64), d.astype(np.float64)
    oc = o - centre
    a = np.einsum("ij,ij->i", d, d)
    b = np.einsum("ij,ij->i", oc, d)
    c0 = np.einsum("ij,ij->i", oc, oc) - radius * radius
    disc = b * b - a * c0
    sq = np.sqrt(np.maximum(disc, 0))
    t0, t1 = (-b - sq) / a, (-b + sq) / a
    t = np.where(t0 > 0, t0, np.where(t1 > 0, t1, np.inf))
    t = np.where(disc >= 0, t, np.inf)
    return t, disc / (a * radius * radius)


def test_sphere_hits_against_the_float64_quadratic(ctx, get_case):
    c = get_case("cornell_spheres")
    ctx.set_scene(c.handle); ctx.set_bvh(c.handle.build_bvh(tie_rank=c.rank)); ctx.set_camera(c.cam)
    ctx.set_options(traversal=api.TRAVERSAL_WIDE, tri_test=api.TRI_REFEXACT)
    sph = [(np.array(s[0], np.float32).astype(np.float64), np.float64(np.float32(s[1]))) for s in c.scene.spheres]
    rng = np.random.default_rng(4)
    pos = c.verts.view(np.float32).reshape(-1, 3)
    n = 200000
    o = rng.uniform(pos.min(0) + 0.02, pos.max(0) - 0.02, size=(n, 3))
    d = rng.normal(size=(n, 3))
    # half of the rays aimed at a sphere's silhouette: centre + r (1 + delta) u with u perpendicular to the line of sight
    k = n // 2
    which = rng.integers(0, len(sph), k)
    cen = np.stack([sph[i][0] for i in which]); rad = np.array([sph[i][1] for i in which])
    los = cen - o[:k]
    u = np.cross(los, rng.normal(size=(k, 3))); u /= np.linalg.norm(u, axis=1, keepdims=True)
    delta = rng.choice([-1e-2, -1e-3, -1e-4, -1e-5, 1e-5, 1e-4, 1e-3, 1e-2, -0.5, -0.9], k) * rng.uniform(0.5, 1.5, k)   # grazing ... central
    d[:k] = cen + u * (rad * (1 + delta))[:, None] - o[:k]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d], axis=1).astype(np.float32)
    inside = np.zeros(n, bool)
    for cc, rr in sph:                                                   # origins inside a sphere: ill-conditioned by construction, left out
        inside |= np.linalg.norm(rays[:, :3].astype(np.float64) - cc, axis=1) < rr * 1.05
    rays = rays[~inside]
    ids, ts = ctx.trace_rays(rays)
    t64 = np.full(len(rays), np.inf); who = np.full(len(rays), -1); margin = np.full(len(rays), np.inf); k1 = np.zeros(len(rays))
    for i, (cc, rr) in enumerate(sph):
        t, m = sphere_roots_f64(rays[:, :3], rays[:, 3:], cc, rr)
        margin = np.minimum(margin, np.abs(m))
        oc2 = ((rays[:, :3].astype(np.float64) - cc) ** 2).sum(1)
        better = t < t64
        t64[better] = t[better]; who[better] = i; k1[better] = (oc2 / np.abs(oc2 - rr * rr))[better]
    gpu_sphere = ids < -1
    gpu_index = ids & 0x7fffffff
    assert gpu_sphere.sum() > 20000
    # (1) where the GPU reports a sphere: it is the sphere the quadratic finds first, at the same distance
    sure = margin > 1e-6
    chk = gpu_sphere & sure
    assert np.array_equal(gpu_index[chk], who[chk])
    err = np.abs(ts[chk].astype(np.float64) - t64[chk]) / (np.abs(t64[chk]) * ULP)
    # The float32 inputs are exact in both evaluations, so the difference is the rounding of the float32 arithmetic times the
    # conditioning of the problem: |c|^2 / ||c|^2 - r^2| (the origin's distance to the surface: cancellation in c.c - r^2) and
    # 1 / sqrt(margin) (grazing incidence: the root's sensitivity to the discriminant).  Well-conditioned hits — origin farther
    # than 0.22 r from the surface, impact parameter below 0.87 r — must agree within 4 ulp; every hit within 4 ulp x conditioning
    # (measured: 3.3 ulp, and 1.7 x conditioning at worst).
    cond = 1 + k1[chk] + 1 / np.sqrt(margin[chk])
    well = (k1[chk] < 3) & (margin[chk] > 0.25)
    print(f"sphere hits: {chk.sum()} checked ({well.sum()} well-conditioned: max error {err[well].max():.2f} ulp); all: max {err.max():.1f} ulp = {(err / cond).max():.2f} x conditioning")
    assert well.sum() > 5000 and err[well].max() <= 4.0
    assert np.all(err <= 4.0 * cond)
    # (2) where the GPU reports a triangle or a miss: no sphere root lies in front of that hit
    other = ~gpu_sphere & sure
    assert np.all(t64[other] >= ts[other].astype(np.float64) * (1 - 1e-5))
    # (3) decisions at the silhouette: outside tangency by more than 1e-6 of the radius -> never a sphere hit nearer than everything else
    miss64 = np.isinf(t64) & sure
    assert not np.any(gpu_sphere & miss64)
    print(f"silhouette rays within 1e-6 of tangency (undecided by construction): {(~sure).sum()} of {len(rays)}")


def test_mirror_reflection_against_the_closed_form(ctx):
    # (the camera sits at (0, 1, 3) looking down -z with slopes up to 0.5: a ray that drops to the floor over a distance L rises to
    # y = 2 over 2 L more, so the emitter hangs 7 to 10 units down the hall)
    FX0, FX1, FZ0, FZ1 = -4.0, 4.0, -6.0, 2.5
    LX0, LX1, LZ0, LZ1 = -1.0, 1.25, -7.0, -4.0
    verts = np.array([[FX0, 0, FZ0], [FX1, 0, FZ0], [FX1, 0, FZ1], [FX0, 0, FZ1],      # mirror floor, y = 0
                      [LX0, 2, LZ0], [LX1, 2, LZ0], [LX1, 2, LZ1], [LX0, 2, LZ1]], np.float32)   # emitter, y = 2, facing down
    tris = np.zeros(4, dtype=api.TRIANGLE)
    tris["v0"], tris["v1"], tris["v2"] = [0, 0, 4, 4], [2, 3, 5, 6], [1, 2, 6, 7]
    tris["norm"][:2] = (0, 1, 0); tris["norm"][2:] = (0, -1, 0)
    tris["mat"] = [0, 0, 1, 1]
    mats = np.zeros(2, dtype=api.MATERIAL)
    mats["albedo"][0] = (0.8, 0.5, 0.25); mats["albedo"][1] = (0.5, 0.5, 0.5); mats["emission"][1] = (3.0, 2.0, 1.0)
    s = api.SceneHandle()
    s.set_arrays(verts.view(api.VEC3).reshape(-1), tris, mats, np.array([2, 3], np.int32), (LX1 - LX0) * (LZ1 - LZ0))
    s.set_specular(0)
    cam = api.make_camera(96, 80, pos=(0.0, 1.0, 3.0), distFromFilm=1.0, focalLength=3.0, radius=0.0)
    ctx.set_scene(s); ctx.set_bvh(s.build_bvh()); ctx.set_camera(cam)
    ctx.set_options(traversal=api.TRAVERSAL_WIDE, tri_test=api.TRI_REFEXACT, integrator=api.INTEGRATOR_PT, max_depth=2, seed=5)
    ctx.reset_accum(); ctx.render(0, 1)
    img = ctx.read_accum(False).astype(np.float64)
    rays = ctx.camera_rays().astype(np.float64)
    o, d = rays[:, :3], rays[:, 3:]
    want = np.zeros_like(img)
    unsure = np.zeros(len(img), bool)

    def in_rect(p, x0, x1, z0, z1, eps):
        inside = (p[:, 0] > x0 + eps) & (p[:, 0] < x1 - eps) & (p[:, 2] > z0 + eps) & (p[:, 2] < z1 - eps)
        near = (np.abs(p[:, 0] - x0) < eps) | (np.abs(p[:, 0] - x1) < eps) | (np.abs(p[:, 2] - z0) < eps) | (np.abs(p[:, 2] - z1) < eps)
        return inside, near & (p[:, 0] > x0 - eps) & (p[:, 0] < x1 + eps) & (p[:, 2] > z0 - eps) & (p[:, 2] < z1 + eps)

    up = d[:, 1] > 0
    with np.errstate(divide="ignore", invalid="ignore"):
        t_up = (2.0 - o[:, 1]) / d[:, 1]
        p_up = o + d * t_up[:, None]
        hit, edge = in_rect(p_up, LX0, LX1, LZ0, LZ1, 1e-3)
        want[up & hit] = (3.0, 2.0, 1.0)                                   # the primary ray meets the emitter: its emission, weight 1
        unsure |= up & edge
        down = d[:, 1] < 0
        t_dn = (0.0 - o[:, 1]) / d[:, 1]
        p_dn = o + d * (t_dn - 0.001)[:, None]                              # the integrator steps back by 0.001 (kernel.cu:431,449)
        on_floor, floor_edge = in_rect(p_dn, FX0, FX1, FZ0, FZ1, 1e-3)
        r = d - 2.0 * d[:, 1:2] * np.array([[0.0, 1.0, 0.0]])               # d - 2 (d.n) n with n = (0, 1, 0)
        t_r = (2.0 - p_dn[:, 1]) / r[:, 1]
        q = p_dn + r * t_r[:, None]
        lit, lit_edge = in_rect(q, LX0, LX1, LZ0, LZ1, 1e-3)
    sel = down & on_floor & lit
    want[sel] = np.array([0.8, 0.5, 0.25]) * np.array([3.0, 2.0, 1.0])      # weight = mirror albedo, then the emitter's emission
    unsure |= down & (floor_edge | (on_floor & lit_edge))
    ok = ~unsure
    assert sel.sum() > 100 and (up & hit).sum() > 20 and ok.mean() > 0.95
    assert np.allclose(img[ok], want[ok], rtol=1e-6, atol=1e-7), np.abs(img[ok] - want[ok]).max()
