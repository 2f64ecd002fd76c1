"""Procedural scenes for the configs of BASELINE.json (the reference ships no models: its .gitignore:1
ignores models/*).  Everything is deterministic numpy; a scene is a material table plus a list of shapes
with ONE material each, which is the granularity of the reference's loadOBJ (modelLoader.h:189).

  cornell_box()        config 1 / 5 base: McGuire-style Cornell box, 32 triangles, framed by the shipped
                       camera pos(0,1,3), distFromFilm 1, focalLength 3 (reference kernel.cu:642-648)
  atrium(n_tris)       config 2 / 3: ~262 144-triangle two-storey colonnaded hall of Sponza proportions
                       (23 x 15 x 37 units, long axis z), open roof closed by a down-facing emissive quad pair
  triangle_soup(n)     config 4: uniform random triangles in [-50,50]^3, edge length log-uniform [0.05,0.5]

`to_obj(scene)` renders a scene to OBJ/MTL text so the same data can go through the OBJ ingest path.
"""
from __future__ import annotations

import dataclasses
from typing import List, Tuple

import numpy as np


@dataclasses.dataclass
class Shape:
    name: str
    positions: np.ndarray  # float32 [nv, 3]
    indices: np.ndarray    # uint32  [nt, 3]
    material: int          # index into Scene.materials


@dataclasses.dataclass
class Scene:
    materials: List[Tuple[str, Tuple[float, float, float], Tuple[float, float, float]]]  # name, Kd, Ke
    shapes: List[Shape]
    camera: dict  # pos, distFromFilm, focalLength, radius (reference camera.h:26-35)
    spheres: list = dataclasses.field(default_factory=list)  # (centre xyz, radius, diffuse rgb, emm rgb): sphere.h:7-13
    specular: list = dataclasses.field(default_factory=list)  # material indices that are perfect mirrors (config 5; not in the reference)

    @property
    def num_triangles(self) -> int:
        return int(sum(s.indices.shape[0] for s in self.shapes))


# --------------------------------------------------------------------------------------------------
# mesh helpers
# --------------------------------------------------------------------------------------------------
def quad_grid(p0, du, dv, nu: int, nv: int):
    """nu x nv quads spanning p0 + s*du + t*dv; geometric normal along cross(du, dv)."""
    p0, du, dv = (np.asarray(a, dtype=np.float64) for a in (p0, du, dv))
    s = np.linspace(0.0, 1.0, nu + 1)
    t = np.linspace(0.0, 1.0, nv + 1)
    S, T = np.meshgrid(s, t, indexing="ij")
    pos = p0[None, None, :] + S[..., None] * du[None, None, :] + T[..., None] * dv[None, None, :]
    pos = pos.reshape(-1, 3).astype(np.float32)
    i, j = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    a = (i * (nv + 1) + j).ravel()
    b = ((i + 1) * (nv + 1) + j).ravel()
    c = ((i + 1) * (nv + 1) + j + 1).ravel()
    d = (i * (nv + 1) + j + 1).ravel()
    tris = np.empty((2 * nu * nv, 3), dtype=np.uint32)
    tris[0::2] = np.stack([a, b, c], axis=1)
    tris[1::2] = np.stack([a, c, d], axis=1)
    return pos, tris


def orient(pos, tris, point, towards: bool):
    """flip triangle windings so that every geometric normal points towards / away from `point`"""
    p = pos.astype(np.float64)
    a, b, c = p[tris[:, 0]], p[tris[:, 1]], p[tris[:, 2]]
    n = np.cross(b - a, c - a)
    d = np.asarray(point, dtype=np.float64)[None, :] - (a + b + c) / 3.0
    s = np.einsum("ij,ij->i", n, d)
    flip = (s < 0) if towards else (s > 0)
    out = tris.copy()
    out[flip] = out[flip][:, ::-1]
    return pos, out


def merge(parts):
    """concatenate (positions, indices) pairs into one indexed mesh"""
    pos, idx, base = [], [], 0
    for p, t in parts:
        pos.append(p)
        idx.append(t.astype(np.int64) + base)
        base += p.shape[0]
    return np.concatenate(pos).astype(np.float32), np.concatenate(idx).astype(np.uint32)


def box_from_top(top, y0=0.0):
    """five faces (top + four sides, outward normals) of a prism whose top quad is `top` (counter-clockwise
    seen from above) extruded down to y0 — the two blocks of the Cornell box"""
    top = np.asarray(top, dtype=np.float64)
    parts = [quad_grid(top[0], top[3] - top[0], top[1] - top[0], 1, 1)]
    for k in range(4):
        a, b = top[k], top[(k + 1) % 4]
        a0, b0 = a.copy(), b.copy()
        a0[1] = y0
        b0[1] = y0
        parts.append(quad_grid(a0, a - a0, b0 - a0, 1, 1))
    pos, tris = merge(parts)
    centre = np.array([top[:, 0].mean(), 0.5 * (top[:, 1].mean() + y0), top[:, 2].mean()])
    return orient(pos, tris, centre, towards=False)


def cylinder(cx, cz, y0, y1, radius, nseg, nrings):
    """open tube, outward normals"""
    ang = np.linspace(0.0, 2.0 * np.pi, nseg + 1)
    ys = np.linspace(y0, y1, nrings + 1)
    A, Y = np.meshgrid(ang, ys, indexing="ij")
    pos = np.stack([cx + radius * np.cos(A), Y, cz + radius * np.sin(A)], axis=-1).reshape(-1, 3).astype(np.float32)
    i, j = np.meshgrid(np.arange(nseg), np.arange(nrings), indexing="ij")
    a = (i * (nrings + 1) + j).ravel()
    b = ((i + 1) * (nrings + 1) + j).ravel()
    c = ((i + 1) * (nrings + 1) + j + 1).ravel()
    d = (i * (nrings + 1) + j + 1).ravel()
    tris = np.empty((2 * nseg * nrings, 3), dtype=np.uint32)
    tris[0::2] = np.stack([a, c, b], axis=1)
    tris[1::2] = np.stack([a, d, c], axis=1)
    return pos, tris


def arch(x, z0, z1, y_spring, thickness, depth, nseg):
    """semicircular arch in the y-z plane between two columns at z0 and z1 (x fixed): soffit strip plus the
    two flat faces of the ring (inner radius r, outer r + thickness), `depth` wide in x"""
    r = 0.5 * abs(z1 - z0)
    zc = 0.5 * (z0 + z1)
    th = np.linspace(0.0, np.pi, nseg + 1)
    parts = []
    xs = (x - 0.5 * depth, x + 0.5 * depth)
    ring_in = np.stack([np.zeros_like(th), y_spring + r * np.sin(th), zc + r * np.cos(th)], axis=-1)
    ring_out = np.stack([np.zeros_like(th), y_spring + (r + thickness) * np.sin(th), zc + (r + thickness) * np.cos(th)], axis=-1)
    # soffit (faces down/inwards)
    pos = np.concatenate([ring_in + [xs[0], 0, 0], ring_in + [xs[1], 0, 0]]).astype(np.float32)
    k = np.arange(nseg)
    t0 = np.stack([k, k + 1, k + nseg + 2], axis=1)
    t1 = np.stack([k, k + nseg + 2, k + nseg + 1], axis=1)
    parts.append(orient(pos, np.concatenate([t0, t1]).astype(np.uint32), (x, y_spring, zc), towards=True))
    # two faces
    for xi in xs:
        pos = np.concatenate([ring_in + [xi, 0, 0], ring_out + [xi, 0, 0]]).astype(np.float32)
        a0 = np.stack([k, k + nseg + 1, k + nseg + 2], axis=1)
        a1 = np.stack([k, k + nseg + 2, k + 1], axis=1)
        tr = np.concatenate([a0, a1]).astype(np.uint32)
        parts.append(orient(pos, tr, (x, y_spring + 0.5 * r, zc), towards=False))
    return merge(parts)


# --------------------------------------------------------------------------------------------------
# scenes
# --------------------------------------------------------------------------------------------------
REFERENCE_CAMERA = dict(pos=(0.0, 1.0, 3.0), distFromFilm=1.0, focalLength=3.0, radius=0.0)  # kernel.cu:642-646


def cornell_box(extra_lights: int = 0, spheres: bool = False, specular: bool = False) -> Scene:
    """McGuire-style Cornell box: 32 triangles, all normals facing the room.
    `extra_lights` > 0 adds that many small emissive triangles under the ceiling (config 5 needs >= 8
    emissive triangles)."""
    white, red, green = (0.725, 0.71, 0.68), (0.63, 0.065, 0.05), (0.14, 0.45, 0.091)
    mats = [("floor", white, (0, 0, 0)), ("ceiling", white, (0, 0, 0)), ("backWall", white, (0, 0, 0)),
            ("rightWall", green, (0, 0, 0)), ("leftWall", red, (0, 0, 0)), ("shortBox", white, (0, 0, 0)),
            ("tallBox", white, (0, 0, 0)), ("light", (0.78, 0.78, 0.78), (17.0, 12.0, 4.0))]
    x0, x1, y1, z0, z1 = -1.01, 1.0, 1.99, -1.04, 0.99
    shapes = []

    def add(name, mesh, m):
        shapes.append(Shape(name, mesh[0], mesh[1], m))

    add("floor", quad_grid((x0, 0, z0), (0, 0, z1 - z0), (x1 - x0, 0, 0), 1, 1), 0)
    add("ceiling", quad_grid((x0, y1, z0), (x1 - x0, 0, 0), (0, 0, z1 - z0), 1, 1), 1)
    add("backWall", quad_grid((x0, 0, z0), (x1 - x0, 0, 0), (0, y1, 0), 1, 1), 2)
    add("rightWall", quad_grid((x1, 0, z0), (0, 0, z1 - z0), (0, y1, 0), 1, 1), 3)
    add("leftWall", quad_grid((x0, 0, z0), (0, y1, 0), (0, 0, z1 - z0), 1, 1), 4)
    add("shortBox", box_from_top([(0.53, 0.6, 0.75), (0.7, 0.6, 0.17), (0.13, 0.6, 0.0), (-0.05, 0.6, 0.57)]), 5)
    add("tallBox", box_from_top([(-0.53, 1.2, 0.09), (0.04, 1.2, -0.09), (-0.14, 1.2, -0.67), (-0.71, 1.2, -0.49)]), 6)
    add("light", quad_grid((-0.24, 1.98, -0.22), (0.47, 0, 0), (0, 0, 0.38), 1, 1), 7)
    if extra_lights > 0:
        rng = np.random.default_rng(5)
        parts = []
        for _ in range(extra_lights):
            cx, cz = rng.uniform(-0.8, 0.8), rng.uniform(-0.8, 0.8)
            p = np.array([(cx, 1.97, cz), (cx + 0.12, 1.97, cz), (cx, 1.97, cz + 0.12)], dtype=np.float32)
            parts.append((p, np.array([[0, 1, 2]], dtype=np.uint32)))
        add("lights2", merge(parts), 7)
    sph = []
    if spheres:   # config 1: "Cornell-box sphere scene (sphere.h)": a diffuse sphere on the short box, an emissive one in the air
        sph = [((0.33, 0.85, 0.37), 0.25, (0.5, 0.75, 0.25), (0.0, 0.0, 0.0)),
               ((-0.6, 1.5, 0.3), 0.125, (0.75, 0.75, 0.75), (4.0, 3.0, 2.0))]
    # config 5 ("mixed diffuse/specular"): the tall box and the back wall become mirrors
    return Scene(mats, shapes, dict(REFERENCE_CAMERA), sph, [6, 2] if specular else [])


ATRIUM_CAMERA = dict(pos=(0.0, 2.2, 17.5), distFromFilm=1.0, focalLength=3.0, radius=0.0)


def atrium_camera() -> dict:
    return dict(ATRIUM_CAMERA)


def atrium(n_tris: int = 262144, seed: int = 1) -> Scene:
    """Two-storey colonnaded hall, 23 x 15 x 37 units (x, y, z), long axis z, camera at the +z end looking
    down -z (the reference camera has no orientation, camera.h:77-97).  Exactly `n_tris` triangles: the
    columns / arches / walls are fixed, the floor grid absorbs the remainder."""
    rng = np.random.default_rng(seed)
    mats = [("floor", (0.55, 0.5, 0.45), (0, 0, 0)), ("wall", (0.7, 0.65, 0.55), (0, 0, 0)),
            ("column", (0.75, 0.72, 0.68), (0, 0, 0)), ("arch", (0.62, 0.58, 0.52), (0, 0, 0)),
            ("gallery", (0.5, 0.42, 0.36), (0, 0, 0)), ("drape", (0.6, 0.14, 0.1), (0, 0, 0)),
            ("skylight", (0.8, 0.8, 0.8), (6.0, 5.5, 5.0))]
    X, Y, Z = 11.5, 15.0, 18.5
    scale = max(0.05, (n_tris / 262144.0) ** 0.5)
    shapes = []
    col_parts, arch_parts = [], []
    col_x, zs = 5.5, np.arange(-16.0, 16.01, 4.0)
    nseg, nrings = max(6, int(round(48 * scale))), max(2, int(round(30 * scale)))
    aseg = max(4, int(round(96 * scale)))
    for storey, (y0, y1) in enumerate(((0.0, 4.5), (7.0, 11.0))):
        for sx in (-col_x, col_x):
            for z in zs:
                col_parts.append(cylinder(sx, float(z), y0, y1, 0.45 - 0.05 * storey, nseg, nrings))
            for za, zb in zip(zs[:-1], zs[1:]):
                arch_parts.append(arch(sx, float(za), float(zb), y1, 0.5, 0.9, aseg))
    shapes.append(Shape("columns", *merge(col_parts), 2))
    shapes.append(Shape("arches", *merge(arch_parts), 3))
    g = max(1, int(round(40 * scale)))
    wall_parts = [
        quad_grid((-X, 0, -Z), (0, Y, 0), (0, 0, 2 * Z), g, 2 * g),          # left wall  (+x normal)
        quad_grid((X, 0, -Z), (0, 0, 2 * Z), (0, Y, 0), 2 * g, g),           # right wall (-x normal)
        quad_grid((-X, 0, -Z), (2 * X, 0, 0), (0, Y, 0), g, g),              # far wall   (+z normal)
        quad_grid((-X, 0, Z), (0, Y, 0), (2 * X, 0, 0), g, g),               # near wall  (-z normal), behind the camera
    ]
    shapes.append(Shape("walls", *merge(wall_parts), 1))
    gal_parts = []
    for sx in (-1.0, 1.0):
        xa, xb = (sx * col_x, sx * X) if sx > 0 else (sx * X, sx * col_x)
        gal_parts.append(quad_grid((xa, 6.5, -Z), (xb - xa, 0, 0), (0, 0, 2 * Z), g // 2 + 1, 2 * g))      # underside (-y)
        gal_parts.append(quad_grid((xa, 7.0, -Z), (0, 0, 2 * Z), (xb - xa, 0, 0), 2 * g, g // 2 + 1))      # top (+y)
        xe = sx * col_x
        gal_parts.append(orient(*quad_grid((xe, 6.5, -Z), (0, 0.5, 0), (0, 0, 2 * Z), 1, 2 * g), (0, 6.75, 0), towards=True))
        # roof over the gallery (-y normal) at y = 13
        gal_parts.append(quad_grid((xa, 13.0, -Z), (xb - xa, 0, 0), (0, 0, 2 * Z), g // 2 + 1, 2 * g))
    shapes.append(Shape("galleries", *merge(gal_parts), 4))
    drape_parts = []
    for z in zs[1:-1:2]:
        w = 1.6 + 0.4 * rng.random()
        drape_parts.append(orient(*quad_grid((-col_x + 0.6, 7.2, float(z) - w / 2), (0, 3.0, 0), (0, 0, w), g // 2 + 1, g // 3 + 1),
                                  (0, 8, 0), towards=True))
        drape_parts.append(orient(*quad_grid((col_x - 0.6, 7.2, float(z) - w / 2), (0, 3.0, 0), (0, 0, w), g // 2 + 1, g // 3 + 1),
                                  (0, 8, 0), towards=True))
    shapes.append(Shape("drapes", *merge(drape_parts), 5))
    # skylight: one quad pair over the nave, facing (0,-1,0) — the reference's PT hard-codes that light normal (kernel.cu:503)
    shapes.append(Shape("skylight", *quad_grid((-col_x, Y - 0.1, -Z + 1), (2 * col_x, 0, 0), (0, 0, 2 * Z - 2), 1, 1), 6))
    # roof strips beside the skylight close the box (-y normal)
    roof = [quad_grid((-X, Y, -Z), (2 * X, 0, 0), (0, 0, 2 * Z), 2, 2)]
    shapes.append(Shape("roof", *merge(roof), 1))
    used = sum(s.indices.shape[0] for s in shapes)
    rem = n_tris - used
    if rem < 4 or rem % 2:
        raise ValueError(f"atrium: fixed geometry already has {used} triangles, cannot reach {n_tris}")
    quads = rem // 2
    nu = max(1, int(np.sqrt(quads * (2 * X) / (2 * Z))))
    nv = max(1, quads // nu)
    k = quads - nu * nv
    floor_parts = [quad_grid((-X, 0, -Z), (0, 0, 2 * Z), (2 * X, 0, 0), nv, nu)]
    if k > 0:  # a thin strip just under the far wall takes the remainder so the count is exact
        floor_parts.append(quad_grid((-X, 0.0, -Z - 0.25), (0, 0, 0.25), (2 * X, 0, 0), 1, k))
    shapes.append(Shape("floor", *merge(floor_parts), 0))
    sc = Scene(mats, shapes, atrium_camera())
    assert sc.num_triangles == n_tris, (sc.num_triangles, n_tris)
    return sc


def triangle_soup(n_tris: int, seed: int = 2, extent: float = 50.0, edge=(0.05, 0.5), cam_z: float = 160.0) -> Scene:
    """centres uniform in [-extent,extent]^3 (50), edge length log-uniform in `edge`
    ([0.05,0.5]), random orientation.  One diffuse material plus a small emissive fraction so paths can
    terminate on lights."""
    rng = np.random.default_rng(seed)
    mats = [("soup", (0.7, 0.7, 0.7), (0, 0, 0)), ("glow", (0.8, 0.8, 0.8), (4.0, 4.0, 4.0))]
    n_light = max(2, n_tris // 512)
    shapes = []
    for name, n, m in (("soup", n_tris - n_light, 0), ("glow", n_light, 1)):
        c = rng.uniform(-extent, extent, size=(n, 1, 3)).astype(np.float32)
        elen = np.exp(rng.uniform(np.log(edge[0]), np.log(edge[1]), size=(n, 1, 1))).astype(np.float32)
        d = rng.normal(size=(n, 3, 3)).astype(np.float32)
        d /= np.linalg.norm(d, axis=2, keepdims=True)
        pos = (c + d * elen * np.float32(0.6)).reshape(-1, 3).astype(np.float32)
        idx = np.arange(3 * n, dtype=np.uint32).reshape(n, 3)
        shapes.append(Shape(name, pos, idx, m))
    return Scene(mats, shapes, dict(pos=(0.0, 0.0, cam_z), distFromFilm=1.0, focalLength=3.0, radius=0.0))


def to_obj(scene: Scene, mtl_name: str = "scene.mtl"):
    """(obj_text, mtl_text): one `o` + `usemtl` block per shape, %.9g so float32 values survive the text"""
    mtl = []
    for name, kd, ke in scene.materials:
        mtl.append(f"newmtl {name}\nKd {kd[0]:.9g} {kd[1]:.9g} {kd[2]:.9g}\nKe {ke[0]:.9g} {ke[1]:.9g} {ke[2]:.9g}\n")
    out = [f"mtllib {mtl_name}\n"]
    base = 0
    for sh in scene.shapes:
        out.append(f"o {sh.name}\n")
        out.append("".join(f"v {p[0]:.9g} {p[1]:.9g} {p[2]:.9g}\n" for p in sh.positions.tolist()))
        out.append(f"usemtl {scene.materials[sh.material][0]}\n")
        idx = sh.indices.astype(np.int64) + base + 1
        out.append("".join(f"f {a} {b} {c}\n" for a, b, c in idx.tolist()))
        base += sh.positions.shape[0]
    return "".join(out), "\n".join(mtl)
