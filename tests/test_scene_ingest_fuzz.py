"""Differential fuzz of scene ingest: OBJ / MTL texts drawn by hypothesis go through the product's loader (csrc/obj_loader.cpp,
csrc/scene.cpp) and through the REAL reference (loadOBJ, modelLoader.h:125-210, with its vendored tiny_obj_loader.cc, compiled into
oracle/_ref by oracle/build_ref.sh); vertex, triangle, material and light arrays and the total light area must be equal byte for byte.
Needs /root/reference at build time, so it runs in the authoring container and skips on the GPU box (fixture `ref`).

What is varied: the spelling of numbers (sign, leading '.', trailing '.', exponents, many digits, leading zeros, values that underflow
or are huge), index spellings (v, v/vt, v/vt/vn, v//vn, negative = relative), polygons (fans), shape splits (g / o / usemtl), white
space (tabs, runs of blanks, trailing blanks, CRLF), comments and blank lines, the order and presence of Kd / Ke / Ka in the MTL."""
import os

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from cudapathtracer_b200 import api


def number(draw):
    kind = draw(st.sampled_from(["int", "dec", "dec", "exp", "lead_dot", "trail_dot", "long", "tiny", "big", "plus", "zeros", "negzero"]))
    sign = draw(st.sampled_from(["", "", "-"]))
    i = draw(st.integers(0, 999))
    f = draw(st.integers(0, 999999))
    if kind == "int":
        return f"{sign}{i}"
    if kind == "dec":
        return f"{sign}{i}.{f}"
    if kind == "exp":
        e = draw(st.integers(-12, 6)); E = draw(st.sampled_from(["e", "E"])); es = draw(st.sampled_from(["", "+"])) if e >= 0 else ""
        return f"{sign}{i % 10}.{f}{E}{es}{e}"
    if kind == "lead_dot":
        return f"{sign}.{f}"
    if kind == "trail_dot":
        return f"{sign}{i}."
    if kind == "long":
        return f"{sign}{i}.{f}{draw(st.integers(0, 10 ** 18))}"
    if kind == "tiny":
        return f"{sign}{i % 10}.{f}e-{draw(st.integers(30, 60))}"
    if kind == "big":
        return f"{sign}{i % 10}.{f}e{draw(st.integers(10, 30))}"
    if kind == "plus":
        return f"+{i}.{f}"
    if kind == "zeros":
        return f"{sign}000{i}.{f}00"
    return "-0"


@st.composite
def obj_files(draw):
    ws = lambda: draw(st.sampled_from([" ", " ", "  ", "\t", " \t "]))   # noqa: E731
    eol = draw(st.sampled_from(["\n", "\r\n"]))
    nv = draw(st.integers(3, 14))
    names = ["m0", "m1", "light"][: draw(st.integers(1, 3))]
    lines = ["mtllib scene.mtl"]
    if draw(st.booleans()):
        lines.append("# a comment")
    for _ in range(nv):
        lines.append("v" + ws() + number(draw) + ws() + number(draw) + ws() + number(draw) + draw(st.sampled_from(["", " ", "\t"])))
        if draw(st.integers(0, 9)) == 0:
            lines.append("")
    nvt, nvn = draw(st.integers(1, 3)), draw(st.integers(1, 3))
    for _ in range(nvt):
        lines.append(f"vt {number(draw)} {number(draw)}")
    for _ in range(nvn):
        lines.append(f"vn {number(draw)} {number(draw)} {number(draw)}")
    lines.append("usemtl" + ws() + draw(st.sampled_from(names)))
    nf = draw(st.integers(1, 10))
    for _ in range(nf):
        split = draw(st.sampled_from(["", "", "", "g", "o", "usemtl", "gusemtl"]))
        if "g" in split[:1]:
            lines.append("g" + ws() + draw(st.sampled_from(["a", "b c", "grp"])))
        if split == "o":
            lines.append("o" + ws() + "obj")
        if "usemtl" in split:
            lines.append("usemtl" + ws() + draw(st.sampled_from(names)))
        k = draw(st.sampled_from([3, 3, 3, 4, 5]))
        fmt = draw(st.sampled_from(["v", "v/t", "v/t/n", "v//n"]))
        neg = draw(st.booleans())
        corners = []
        for _ in range(k):
            vi = draw(st.integers(1, nv))
            vtok = str(vi - nv - 1) if neg else str(vi)
            ti = draw(st.integers(1, nvt)); ni = draw(st.integers(1, nvn))
            ttok = str(ti - nvt - 1) if neg else str(ti)
            ntok = str(ni - nvn - 1) if neg else str(ni)
            corners.append({"v": vtok, "v/t": f"{vtok}/{ttok}", "v/t/n": f"{vtok}/{ttok}/{ntok}", "v//n": f"{vtok}//{ntok}"}[fmt])
        lines.append("f" + ws() + ws().join(corners) + draw(st.sampled_from(["", " "])))
    obj = eol.join(lines) + eol
    mtl_lines = []
    for n in names:
        mtl_lines.append("newmtl " + n)
        props = [("Kd", f"{number(draw)} {number(draw)} {number(draw)}")]
        if n == "light" or draw(st.booleans()):
            ke = "4 3.5 3" if n == "light" else draw(st.sampled_from(["0 0 0", "0 1 0", "0.0 0.0 0.0"]))
            props.append(("Ke", ke))
        if draw(st.booleans()):
            props.append(("Ka", "1 1 1"))
        for key, val in draw(st.permutations(props)):
            mtl_lines.append(draw(st.sampled_from(["", "  ", "\t"])) + key + " " + val)
    mtl = eol.join(mtl_lines) + eol
    return obj, mtl


@settings(max_examples=int(os.environ.get("CUPT_FUZZ_EXAMPLES", "250")), deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow], derandomize=True)
@given(obj_files(), st.sampled_from([((0.0, 0.0, 0.0), 1.0, False), ((0.35, -2.0, 1e3), 0.75, True)]))
def test_generated_obj_files_load_like_the_reference(lib, ref, tmp_path_factory, files, xform):
    obj, mtl = files
    origin, scale, flip = xform
    ref.reset()
    ref.load_obj_text(obj, mtl, origin=origin, scale=scale, flip=flip)
    rv, rt, rm, rl, ra = ref.arrays()
    d = tmp_path_factory.mktemp("fuzz") / "models"
    d.mkdir()
    (d / "scene.obj").write_bytes(obj.encode())
    (d / "scene.mtl").write_bytes(mtl.encode())
    s = api.SceneHandle()
    s.load_obj(str(d / "scene.obj"), str(d) + "/", origin=origin, scale=scale, flip=flip)
    v, t, m, l = s.arrays()
    assert len(t) == len(rt) and len(v) == len(rv), (obj, mtl)
    assert v.tobytes() == rv.tobytes(), (obj, v, rv)
    assert t.tobytes() == rt.tobytes(), (obj, t, rt)
    assert m.tobytes() == rm.tobytes(), (mtl, m, rm)
    assert l.tobytes() == rl.tobytes() and (s.total_light_area == ra or (np.isnan(ra) and np.isnan(s.total_light_area))), (obj, mtl, l, rl, s.total_light_area, ra)
    s.close()
