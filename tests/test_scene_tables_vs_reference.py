"""Pins the host scene layer (rustraytracer_b200/csrc/host_scenes.cpp) against the reference's own source text.

The GPU path and the oracle are fed the SAME tables from HostScene.tables(), so a transcription slip against scenes.rs would be invisible to every
parity test.  Here the numeric literals of each scene function are read from /root/reference/src/scenes.rs at test time
and every constant in our tables — rect extents, sphere centres and radii, texture colours, checker frequencies, material scalars, light colours,
transform entries — must be one of those literals or a product / quotient of at most three of them (`Vector3::new(0.1514, 0.0139,
0.3765).scale(0.56 / 0.3765)`), and the numbers of shapes / textures / materials / lights must equal the numbers of constructor calls in the text.
Runs only where the reference checkout exists (this container); the GPU box does not have it."""
import os
import re

import numpy as np
import pytest

from rustraytracer_b200 import _native as N, scenes

SCENES_RS = "/root/reference/src/scenes.rs"
pytestmark = pytest.mark.skipif(not os.path.exists(SCENES_RS), reason="reference checkout not present")


def fn_body(text, name):
    m = re.search(r"\n(?:pub )?fn " + name + r"\(", text)
    assert m, name
    nxt = re.search(r"\n(?:pub )?fn \w+\(", text[m.end():])
    body = text[m.start(): m.end() + (nxt.start() if nxt else len(text))]
    return re.sub(r"//[^\n]*", "", body)            # constants in comments do not count


def literals(body):
    vals = set()
    for tok in re.findall(r"(?<![\w.])-?\d+\.?\d*(?:e-?\d+)?(?:f64|usize|u32)?", body):
        tok = re.sub(r"(f64|usize|u32)$", "", tok)
        try:
            vals.add(float(tok))
        except ValueError:
            pass
    return vals


def derivable(lits):
    """literals and the products / quotients of up to three of them (and 1 - literal: `1. - 0.718` in rosegold_metal)"""
    arr = np.array(sorted(lits | {1.0, 2.0, 0.5, 256.0, 255.0}))
    two = np.concatenate([np.outer(arr, arr).ravel(), (arr[:, None] / arr[None, arr != 0]).ravel()])
    small = arr[np.abs(arr) <= 1e4]
    three = np.concatenate([np.outer(two, small).ravel(), (two[:, None] / small[None, small != 0]).ravel()])
    allv = np.concatenate([arr, -arr, two, three, 1.0 - arr])
    return allv[np.isfinite(allv)]


def unroll_loops(body):
    """`for i in a..b { ... }` blocks repeated b - a times, so that constructor calls inside them count once per iteration"""
    out, pos = "", 0
    for m in re.finditer(r"for \w+ in (\d+)\.\.(\d+) \{", body):
        if m.start() < pos:
            continue
        depth, j = 1, m.end()
        while depth and j < len(body):
            depth += {"{": 1, "}": -1}.get(body[j], 0); j += 1
        out += body[pos:m.start()] + body[m.end():j - 1] * (int(m.group(2)) - int(m.group(1)))
        pos = j
    return out + body[pos:]


def extras(name, body):
    """values a scene function computes with more than products of literals: the formula is restated here, its numbers are read from the text"""
    out = []
    if name == "sphere_roughness":
        space = float(re.search(r"let space: f64 = ([\d.]+);", body).group(1))
        a = re.search(r"-space \* ([\d.]+) \+ space \* i as f64", body).group(1)
        b = re.search(r"\(\(i - 1\) as f64\) / ([\d.]+) \+ SMALL", body).group(1)
        lo, hi = (int(x) for x in re.search(r"for i in (\d+)\.\.(\d+)", body).groups())
        for i in range(lo, hi):
            out += [-space * float(a) + space * i, (i - 1) / float(b) + 0.001]          # SMALL = 0.001 (consts.rs:32)
    return out


def check_constants(name, consts, raw):
    srt = np.sort(raw)
    for what, c in consts:
        if c == 0.0 or abs(c) == 1.0:
            continue
        i = np.searchsorted(srt, c)
        near = [srt[j] for j in (i - 1, i) if 0 <= j < len(srt)]
        assert any(abs(c - x) <= 1e-12 * max(1.0, abs(c)) for x in near), f"{name}: {what} = {c!r} is not derivable from the literals of scenes.rs"


CASES = [
    ("sphere_roughness", {}, ["sphere_roughness"]),
    ("cornell_box", {}, ["cornell_box"]),
    ("cornell_box_statue", {"n_tris": 2000}, ["cornell_box_statue"]),
    ("plastic_dragon", {}, ["plastic_dragon"]),
    ("two_dragons", {}, ["two_dragons"]),
    ("material_hdr", {"variant": 0}, ["material_hdr", "smooth_plastic"]),
    ("material_hdr", {"variant": 1}, ["material_hdr", "rosegold_metal"]),
    ("material_hdr", {"variant": 2}, ["material_hdr", "mirror"]),
    ("material_hdr", {"variant": 3}, ["material_hdr", "glass"]),
    ("material_hdr", {"variant": 4}, ["material_hdr", "disney_mat"]),
    ("teapot_hdr", {}, ["teapot_hdr"]),
]


@pytest.mark.parametrize("name,kw,fns", CASES)
def test_scene_constants_come_from_scenes_rs(name, kw, fns):
    text = open(SCENES_RS).read()
    body = "\n".join(fn_body(text, f) for f in fns)
    raw = np.concatenate([derivable(literals(body)), np.array(extras(name, body), dtype=np.float64)])
    hs = scenes.HostScene(name, 64, 64, **kw)
    t = hs.tables()
    consts = []
    shapes = t["prims"][t["prims"]["type"] != N.PRIM_TRIANGLE]
    for i, p in enumerate(shapes):
        nv = 4 if p["type"] == N.PRIM_SPHERE else 5
        consts += [(f"shape {i} v[{k}]", float(p["v"][k])) for k in range(nv)]
    for i, x in enumerate(t["textures"]):
        if x["type"] == N.TEX_SOLID:
            consts += [(f"texture {i} colour[{k}]", float(x["color"][k])) for k in range(3)]
        if x["type"] == N.TEX_CHECKERED:
            consts.append((f"texture {i} frequency", float(x["frequency"])))
    for i, m in enumerate(t["materials"]):
        n_f = {N.MAT_MATTE: 1, N.MAT_PLASTIC: 1, N.MAT_GLASS: 3}.get(int(m["type"]), 0)
        consts += [(f"material {i} f[{k}]", float(m["f"][k])) for k in range(n_f)]
    for i, l in enumerate(t["lights"]):
        if l["type"] == N.LIGHT_DIFFUSE:
            consts += [(f"light {i} colour[{k}]", float(l["color"][k])) for k in range(3)]
    if name not in ("cornell_box", "cornell_box_statue"):          # those transforms are Ry(15 deg) / Ry(-18 deg) / Rz(pi) composites: trig of literals
        for i, m in enumerate(np.asarray(t["transforms"]).reshape(-1, 16)):
            consts += [(f"transform {i}[{k}]", float(m[k])) for k in range(12) if abs(m[k]) > 1e-6]
    check_constants(name + str(kw), consts, raw)
    # structure: constructor calls in the text == entries in our tables
    body = unroll_loops(body)
    n_shape_calls = len(re.findall(r"Primitive::new_(?:xy_rect|xz_rect|yz_rect|sphere)\w*\(", body))
    n_cubes = len(re.findall(r"Cube::new\w*\(|make_cube\w*\(|new_cube\w*\(", body))
    assert n_cubes or len(shapes) == n_shape_calls, (name, len(shapes), n_shape_calls, n_cubes)
    assert len(t["textures"]) == len(re.findall(r"Texture::new_\w+\(", body)), name
    assert len(t["materials"]) == len(re.findall(r"Material::make_\w+\(", body)), name
    assert len(t["lights"]) == len(re.findall(r"Light::make_\w+\(", body)), name
    assert re.search(r"Camera::new(?:_motion_blur)?\(", body), name
