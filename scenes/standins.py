"""Deterministic stand-ins for the assets the reference checkout lacks
(.MISSING_LARGE_BLOBS), plus the per-scene build
steps that combine them with the shipped files and export the archive.

Every generator is seeded (numpy default_rng(20261104)) or closed-form, so the
CUDA path and the reference CPU renderer always see the same bytes.
"""
from __future__ import annotations

import json
import os
import struct

import numpy as np

from make_scenes import REF_TESTS, WORK, export, link_tree

SEED = 20261104


# --------------------------------------------------------------------------------------
# asset writers
# --------------------------------------------------------------------------------------
def write_obj_quad_sphere(path, steps_u=64, steps_v=32, radius=0.075, center=(0, 0.075, 0)):
    """Quad UV-sphere with texcoords; pole rows are triangles. Same extents as sphere.ply."""
    lines = []
    for j in range(steps_v + 1):
        th = np.pi * j / steps_v
        for i in range(steps_u + 1):
            ph = 2 * np.pi * i / steps_u
            p = (center[0] + radius * np.cos(ph) * np.sin(th), center[1] + radius * np.cos(th),
                 center[2] + radius * np.sin(ph) * np.sin(th))
            lines.append("v %.9g %.9g %.9g" % p)
    for j in range(steps_v + 1):
        for i in range(steps_u + 1):
            lines.append("vt %.9g %.9g" % (i / steps_u, 1 - j / steps_v))
    row = steps_u + 1
    for j in range(steps_v):
        for i in range(steps_u):
            a, b = j * row + i + 1, j * row + i + 2
            c, d = (j + 1) * row + i + 2, (j + 1) * row + i + 1
            if j == 0:
                idx = (a, c, d)
            elif j == steps_v - 1:
                idx = (a, b, d)
            else:
                idx = (a, b, c, d)
            lines.append("f " + " ".join(f"{k}/{k}" for k in idx))
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


def write_obj_lathe(path, profile, steps_u, center=(0, 0, 0), handle=None):
    """Quad surface of revolution of `profile` [(radius, height), ...] (first and last radius 0:
    pole fans of triangles), with texcoords; optionally a quad torus `handle` = (centre, major
    radius, minor radius, steps) in the xy plane. The teapot stand-in."""
    lines, faces = [], []
    nv = len(profile)
    for j, (r, y) in enumerate(profile):
        for i in range(steps_u + 1):
            ph = 2 * np.pi * i / steps_u
            lines.append("v %.9g %.9g %.9g" % (center[0] + r * np.cos(ph), center[1] + y,
                                              center[2] + r * np.sin(ph)))
    vts = [(i / steps_u, 1 - j / (nv - 1)) for j in range(nv) for i in range(steps_u + 1)]
    row = steps_u + 1
    for j in range(nv - 1):
        for i in range(steps_u):
            a, b = j * row + i + 1, j * row + i + 2
            c, d = (j + 1) * row + i + 2, (j + 1) * row + i + 1
            if profile[j][0] == 0:
                faces.append((a, c, d))
            elif profile[j + 1][0] == 0:
                faces.append((a, b, d))
            else:
                faces.append((a, b, c, d))
    if handle is not None:
        hc, major, minor, hs = handle
        base = len(vts)
        for j in range(hs + 1):
            th = 2 * np.pi * j / hs
            for i in range(hs + 1):
                ph = 2 * np.pi * i / hs
                rr = major + minor * np.cos(ph)
                lines.append("v %.9g %.9g %.9g" % (center[0] + hc[0] + rr * np.cos(th),
                                                  center[1] + hc[1] + rr * np.sin(th),
                                                  center[2] + hc[2] + minor * np.sin(ph)))
                vts.append((j / hs, i / hs))
        for j in range(hs):
            for i in range(hs):
                a = base + j * (hs + 1) + i + 1
                faces.append((a, a + 1, a + hs + 2, a + hs + 1))
    lines += ["vt %.9g %.9g" % t for t in vts]
    lines += ["f " + " ".join(f"{k}/{k}" for k in f) for f in faces]
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")


TEAPOT_PROFILE = [(0.0, 0.0), (0.045, 0.0), (0.062, 0.012), (0.072, 0.035), (0.068, 0.060),
                  (0.052, 0.080), (0.040, 0.088), (0.043, 0.092), (0.030, 0.100), (0.012, 0.106),
                  (0.010, 0.114), (0.016, 0.120), (0.0, 0.126)]


def refine_profile(profile, times):
    """Midpoint refinement of a lathe profile (keeps the poles)."""
    for _ in range(times):
        out = []
        for a, b in zip(profile[:-1], profile[1:]):
            out += [a, ((a[0] + b[0]) / 2, (a[1] + b[1]) / 2)]
        profile = out + [profile[-1]]
    return profile


def write_normal_png(path, size=1024):
    """Tangent-space normal map (byte, linear): gradients of tiling value noise."""
    from PIL import Image

    rng = np.random.default_rng(SEED + 2)
    h = sum(value_noise(size, 16 * 2 ** o, rng) / 2 ** o for o in range(3))
    gx = (np.roll(h, -1, axis=1) - np.roll(h, 1, axis=1)) * size / 64.0
    gy = (np.roll(h, -1, axis=0) - np.roll(h, 1, axis=0)) * size / 64.0
    n = np.stack([-gx, -gy, np.ones_like(h)], 2)
    n /= np.linalg.norm(n, axis=2, keepdims=True)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    Image.fromarray(((n * 0.5 + 0.5) * 255 + 0.5).astype(np.uint8)).save(path)


def icosphere(subdivisions):
    """Unit icosphere: (positions [n,3] float64, triangles [m,3] int32)."""
    t = (1 + 5 ** 0.5) / 2
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t],
                  [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], float)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4],
                  [11, 10, 2], [10, 7, 6], [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8],
                  [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], np.int64)
    for _ in range(subdivisions):
        edges = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        key = np.sort(edges, axis=1)
        uniq, inv = np.unique(key, axis=0, return_inverse=True)
        mid = v[uniq[:, 0]] + v[uniq[:, 1]]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        base = len(v)
        v = np.concatenate([v, mid])
        n = len(f)
        m01, m12, m20 = base + inv[:n], base + inv[n:2 * n], base + inv[2 * n:]
        f = np.concatenate([np.stack([f[:, 0], m01, m20], 1), np.stack([f[:, 1], m12, m01], 1),
                            np.stack([f[:, 2], m20, m12], 1), np.stack([m01, m12, m20], 1)])
    return v, f.astype(np.int32)


def displaced_head(subdivisions, center, radius):
    """Noise-displaced icosphere with vertex normals and spherical texcoords."""
    v, f = icosphere(subdivisions)
    rng = np.random.default_rng(SEED)
    disp = np.zeros(len(v))
    for octave in range(5):
        freq = 2.0 * 2 ** octave
        dirs = rng.normal(size=(4, 3))
        phase = rng.uniform(0, 2 * np.pi, size=4)
        for d, ph in zip(dirs, phase):
            disp += (0.08 / 2 ** octave) * np.sin(freq * (v @ d) + ph)
    p = v * (1 + disp)[:, None]
    fn = np.cross(p[f[:, 1]] - p[f[:, 0]], p[f[:, 2]] - p[f[:, 0]])
    n = np.zeros_like(p)
    for k in range(3):
        np.add.at(n, f[:, k], fn)
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20)
    uv = np.stack([np.arctan2(v[:, 2], v[:, 0]) / (2 * np.pi) + 0.5,
                   np.arccos(np.clip(v[:, 1], -1, 1)) / np.pi], 1)
    pos = p * radius + np.asarray(center)[None, :]
    return pos.astype(np.float32), n.astype(np.float32), uv.astype(np.float32), f


def write_ply(path, pos, nrm, uv, tris):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    header = ("ply\nformat binary_little_endian 1.0\n"
              f"element vertex {len(pos)}\n"
              "property float x\nproperty float y\nproperty float z\n"
              "property float nx\nproperty float ny\nproperty float nz\n"
              "property float u\nproperty float v\n"
              f"element face {len(tris)}\n"
              "property list uchar int vertex_indices\nend_header\n")
    verts = np.concatenate([pos, nrm, uv], axis=1).astype("<f4")
    faces = np.zeros(len(tris), dtype=[("n", "u1"), ("i", "<i4", 3)])
    faces["n"], faces["i"] = 3, tris
    with open(path, "wb") as f:
        f.write(header.encode())
        f.write(verts.tobytes())
        f.write(faces.tobytes())


def value_noise(size, cells, rng):
    """Smooth tiling value noise in [0,1], shape [size, size]."""
    g = rng.uniform(size=(cells, cells))
    x = np.arange(size) * cells / size
    i0 = np.floor(x).astype(int) % cells
    i1 = (i0 + 1) % cells
    t = x - np.floor(x)
    t = t * t * (3 - 2 * t)
    rows = g[i0][:, i0] * (1 - t)[None, :] + g[i0][:, i1] * t[None, :]
    rows1 = g[i1][:, i0] * (1 - t)[None, :] + g[i1][:, i1] * t[None, :]
    return rows * (1 - t)[:, None] + rows1 * t[:, None]


def write_noise_png(path, size=2048):
    from PIL import Image

    rng = np.random.default_rng(SEED + 1)
    img = np.zeros((size, size, 3))
    for c in range(3):
        acc = sum(value_noise(size, 8 * 2 ** o, rng) / 2 ** o for o in range(4)) / 1.875
        img[..., c] = 0.35 + 0.6 * acc
    os.makedirs(os.path.dirname(path), exist_ok=True)
    Image.fromarray((np.clip(img, 0, 1) * 255 + 0.5).astype(np.uint8)).save(path)


def write_hdr(path, rgb):
    """Uncompressed Radiance RGBE, rgb: [h, w, 3] float."""
    h, w, _ = rgb.shape
    m = np.max(rgb, axis=2)
    e = np.zeros_like(m, dtype=np.int32)
    nz = m > 1e-32
    mant, ex = np.frexp(m[nz])
    scale = np.zeros_like(m)
    scale[nz] = mant * 256.0 / m[nz]
    e[nz] = ex + 128
    out = np.zeros((h, w, 4), dtype=np.uint8)
    out[..., :3] = np.clip(rgb * scale[..., None], 0, 255).astype(np.uint8)
    out[..., 3] = np.clip(e, 0, 255).astype(np.uint8)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "wb") as f:
        f.write(b"#?RADIANCE\nFORMAT=32-bit_rle_rgbe\n\n")
        f.write(f"-Y {h} +X {w}\n".encode())
        f.write(out.tobytes())


def write_sky_hdr(path, width=1000, height=500):
    th = (np.arange(height) + 0.5) / height * np.pi
    ph = (np.arange(width) + 0.5) / width * 2 * np.pi
    up = np.cos(th)[:, None] * np.ones(width)[None, :]
    sky = np.stack([0.35 + 0.25 * up, 0.45 + 0.3 * up, 0.6 + 0.4 * up], 2)
    sky = np.where(up[..., None] < 0, 0.15, sky)
    d = np.stack([np.cos(ph)[None, :] * np.sin(th)[:, None], up,
                  np.sin(ph)[None, :] * np.sin(th)[:, None]], 2)
    sun = np.array([0.5, 0.6, 0.62])
    sun /= np.linalg.norm(sun)
    c = np.clip(d @ sun, 0, 1)
    sky = sky + 40.0 * (c ** 400)[..., None] * np.array([1.0, 0.9, 0.7])
    write_hdr(path, sky)


def write_sdf_grid(path, n, res, field):
    """SDFGen text grid (yocto_sceneio.cpp:913-941): 'W H D', origin line, res, values x-fastest."""
    idx = (np.arange(n) + 0.0) * res
    z, y, x = np.meshgrid(idx, idx, idx, indexing="ij")
    values = field(x, y, z).astype(np.float32).reshape(n * n, n)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as f:
        f.write(f"{n} {n} {n}\n0 0 0\n{res:.9g}\n")
        np.savetxt(f, values, fmt="%.9g")


# --------------------------------------------------------------------------------------
# distance fields of the stand-in grids
# --------------------------------------------------------------------------------------
def _sphere(x, y, z, c, r):
    return np.sqrt((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) - r


def _capsule(x, y, z, a, b, r):
    a, b = np.asarray(a, float), np.asarray(b, float)
    ba = b - a
    px, py, pz = x - a[0], y - a[1], z - a[2]
    h = np.clip((px * ba[0] + py * ba[1] + pz * ba[2]) / (ba @ ba), 0, 1)
    return np.sqrt((px - ba[0] * h) ** 2 + (py - ba[1] * h) ** 2 + (pz - ba[2] * h) ** 2) - r


def bunny_field(x, y, z):
    """Blobby 'bunny' in a 0.2 m cube: three spheres, two ears, slight ripple."""
    d = _sphere(x, y, z, (0.10, 0.060, 0.10), 0.055)
    d = np.minimum(d, _sphere(x, y, z, (0.10, 0.125, 0.125), 0.034))
    d = np.minimum(d, _sphere(x, y, z, (0.10, 0.055, 0.045), 0.020))
    d = np.minimum(d, _capsule(x, y, z, (0.085, 0.15, 0.12), (0.075, 0.185, 0.105), 0.009))
    d = np.minimum(d, _capsule(x, y, z, (0.115, 0.15, 0.12), (0.125, 0.185, 0.105), 0.009))
    ripple = 0.0015 * np.sin(140 * x) * np.sin(150 * y) * np.sin(130 * z)
    return 0.85 * (d + ripple)


def sackboy_field(x, y, z):
    """Capsule figure in a 200 mm cube (the instance scales by 0.001)."""
    d = _sphere(x, y, z, (100, 150, 100), 30)
    d = np.minimum(d, _capsule(x, y, z, (100, 70, 100), (100, 110, 100), 24))
    d = np.minimum(d, _capsule(x, y, z, (76, 108, 100), (45, 80, 104), 9))
    d = np.minimum(d, _capsule(x, y, z, (124, 108, 100), (155, 80, 104), 9))
    d = np.minimum(d, _capsule(x, y, z, (88, 52, 100), (82, 14, 102), 11))
    d = np.minimum(d, _capsule(x, y, z, (112, 52, 100), (118, 14, 102), 11))
    return 0.9 * d


def write_standin_grids(scene_dir, n=128):
    bunny = os.path.join(scene_dir, "sdfs", "bunny.sdf")
    sack = os.path.join(scene_dir, "sdfs", "sackboy.sdf")
    if not os.path.exists(bunny):
        write_sdf_grid(bunny, n, 0.2 / n, bunny_field)
    if not os.path.exists(sack):
        write_sdf_grid(sack, n, 200.0 / n, sackboy_field)


# --------------------------------------------------------------------------------------
# per-scene builds
# --------------------------------------------------------------------------------------
def _work_dir(name):
    src = os.path.join(REF_TESTS, name)
    dst = os.path.join(WORK, name)
    link_tree(src, dst)
    return dst


def build_01_surface(force=False):
    d = _work_dir("01_surface")
    for sub in ("shapes", "subdivs"):
        p = os.path.join(d, sub, "sphere-displaced.obj")
        if not os.path.exists(p):
            write_obj_quad_sphere(p)
    return export(d, "surface.json", "01_surface", force)


def build_02_rollingteapot(force=False):
    """tests/02_rollingteapot: the four teapot OBJs are missing; a lathe teapot with a torus
    handle stands in (a dense one as the plain shape, a coarse cage for the Catmull-Clark
    subdiv: `subdivisions 1, smooth`)."""
    d = _work_dir("02_rollingteapot")
    dense = os.path.join(d, "shapes", "rollingteapot-shape.obj")
    if not os.path.exists(dense):
        write_obj_lathe(dense, refine_profile(TEAPOT_PROFILE, 2), 96,
                        handle=((0.088, 0.06, 0.0), 0.028, 0.006, 24))
    for sub in ("shapes", "subdivs"):
        cage = os.path.join(d, sub, "rollingteapot-subdiv.obj")
        if not os.path.exists(cage):
            write_obj_lathe(cage, TEAPOT_PROFILE, 24, handle=((0.088, 0.06, 0.0), 0.028, 0.007, 8))
    # the same scene with its subdivision surface left to the device tessellation
    export(d, "rollingteapot.json", "02_rollingteapot_subdiv", force, keep_subdivs=True)
    return export(d, "rollingteapot.json", "02_rollingteapot", force)


def build_04_head1(force=False, subdivisions=None):
    """tests/04_head1: the 05 stand-in mesh, a noise colour texture and a noise normal map;
    texture3.hdr (two environments share it) is shipped."""
    d = _work_dir("04_head1")
    cam = json.load(open(os.path.join(d, "head1.json")))["cameras"][0]
    fr = np.array(cam["frame"], float).reshape(4, 3)
    center = fr[3] - fr[2] * cam["focus"]
    half = cam["focus"] * 0.036 / 2 / cam.get("lens", 0.05)
    radius = 0.6 * half / 1.2
    # 7 subdivisions (327 680 triangles): 04 is not a BASELINE configuration, and the snapshot
    # that travels to the GPU box is capped at 512 MiB
    subdivisions = subdivisions or int(os.environ.get("VPT_HEAD1_SUBDIV", "7"))
    ply = os.path.join(d, "shapes", "shape1.ply")
    if not os.path.exists(ply):
        write_ply(ply, *displaced_head(subdivisions, center, radius))
    png = os.path.join(d, "textures", "texture1.png")
    if not os.path.exists(png):
        write_noise_png(png, size=1024)
    nrm = os.path.join(d, "textures", "texture2.png")
    if not os.path.exists(nrm):
        write_normal_png(nrm)
    return export(d, "head1.json", "04_head1", force)


def build_05_head1ss(force=False, subdivisions=None):
    d = _work_dir("05_head1ss")
    cam = json.load(open(os.path.join(d, "head1ss.json")))["cameras"][0]
    fr = np.array(cam["frame"], float).reshape(4, 3)
    center = fr[3] - fr[2] * cam["focus"]  # the camera looks down -z of its frame
    half = cam["focus"] * 0.036 / 2 / cam["lens"]  # half film extent at the focus plane
    radius = 0.6 * half / 1.2
    subdivisions = subdivisions or int(os.environ.get("VPT_HEAD_SUBDIV", "8"))
    ply = os.path.join(d, "shapes", "shape1.ply")
    if not os.path.exists(ply):
        write_ply(ply, *displaced_head(subdivisions, center, radius))
    png = os.path.join(d, "textures", "texture1.png")
    if not os.path.exists(png):
        write_noise_png(png)
    hdr = os.path.join(d, "textures", "texture2.hdr")
    if not os.path.exists(hdr):
        write_sky_hdr(hdr)
    return export(d, "head1ss.json", "05_head1ss", force)


def build_06_gridsdf(force=False):
    d = _work_dir("06_gridsdf")
    write_standin_grids(d)
    return export(d, "gridsdf.json", "06_gridsdf", force)


def build_07_sdfunction(force=False):
    d = _work_dir("07_sdfunction")
    write_standin_grids(d)
    return export(d, "sdfunction.json", "07_sdfunction", force)
