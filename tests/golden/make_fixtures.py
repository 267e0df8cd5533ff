"""Builds the two committed fixture scenes and their golden oracle outputs.

  mixed.vptscene      every material type, triangles and quads (incl. a quad that
                      degenerates to a triangle), byte sRGB / byte linear / float
                      textures, a normal map, vertex colours, opacity < 1, a
                      non-rigid instance frame, an area light and a textured
                      environment; four cameras (pinhole, thin lens, orthographic, portrait)
  sdf_small.vptscene  one 16^3 voxel-grid SDF instance plus one analytic SDF of every type

The scenes are written from numpy alone (seeded), so they can be rebuilt
anywhere. The golden outputs are produced by oracle/_ref/ref_render — the
UNMODIFIED reference CPU path — and therefore only where /root/reference was
available to build it:

    python tests/golden/make_fixtures.py            # scenes + goldens
    python tests/golden/make_fixtures.py --scenes   # scenes only
"""
from __future__ import annotations

import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

MATTE, GLOSSY, REFLECTIVE, TRANSPARENT, REFRACTIVE, SUBSURFACE, VOLUMETRIC, GLTFPBR = range(8)
SDF_BBOX, SDF_BOX, SDF_CAPPED_CONE, SDF_PLANE, SDF_SPHERE, SDF_TORUS = range(6)

# (scene, shader, samples, resolution, bounces) rendered into tests/golden/*.vptarch
GOLDEN_RENDERS = [
    ("mixed", "color", 1, 64, 4), ("mixed", "texcoord", 1, 64, 4), ("mixed", "normal", 1, 64, 4),
    ("mixed", "eyelight", 2, 64, 4), ("mixed", "naive", 2, 64, 4), ("mixed", "pathtrace", 2, 64, 4),
    ("mixed", "volpathtrace", 2, 64, 8),
    ("sdf_small", "implicit_normal", 1, 64, 4), ("sdf_small", "implicit", 2, 64, 4),
    ("skewed", "color", 1, 64, 4), ("skewed", "pathtrace", 2, 64, 4),
]


def f32(x):
    return np.asarray(x, dtype=np.float32)


def normalize(v):
    v = np.asarray(v, dtype=np.float64)
    return v / np.linalg.norm(v)


def frame(o=(0, 0, 0), scale=(1, 1, 1), rot_y=0.0, rot_x=0.0):
    cy, sy, cx, sx = np.cos(rot_y), np.sin(rot_y), np.cos(rot_x), np.sin(rot_x)
    ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    m = ry @ rx @ np.diag(scale)
    # 12 floats: x axis, y axis, z axis (columns), origin
    return f32(np.concatenate([m[:, 0], m[:, 1], m[:, 2], np.asarray(o, dtype=np.float64)]))


def lookat(eye, center, up=(0, 1, 0)):
    z = normalize(np.subtract(eye, center))
    x = normalize(np.cross(up, z))
    y = normalize(np.cross(z, x))
    return f32(np.concatenate([x, y, z, np.asarray(eye, dtype=np.float64)]))


def camera(eye, center, lens=0.05, film=0.036, aspect=1.5, aperture=0.0, orthographic=False):
    focus = float(np.linalg.norm(np.subtract(eye, center)))
    return np.concatenate([lookat(eye, center),
                           f32([float(orthographic), lens, film, aspect, focus, aperture])])


def material(type_, color=(0, 0, 0), emission=(0, 0, 0), roughness=0.0, metallic=0.0, ior=1.5,
             scattering=(0, 0, 0), scanisotropy=0.0, trdepth=0.01, opacity=1.0, emission_tex=-1,
             color_tex=-1, roughness_tex=-1, scattering_tex=-1, normal_tex=-1):
    f = f32([*emission, *color, roughness, metallic, ior, *scattering, scanisotropy, trdepth,
             opacity, 0])
    i = np.asarray([type_, emission_tex, color_tex, roughness_tex, scattering_tex, normal_tex],
                   dtype=np.int32)
    return f, i


def quad_shape(size=1.0, uvscale=1.0):
    p = f32([[-size, 0, size], [size, 0, size], [size, 0, -size], [-size, 0, -size]])
    n = f32([[0, 1, 0]] * 4)
    t = f32([[0, 0], [uvscale, 0], [uvscale, uvscale], [0, uvscale]])
    return {"positions": p, "normals": n, "texcoords": t, "quads": np.asarray([[0, 1, 2, 3]], np.int32)}


def icosphere(subdiv, radius=1.0, with_normals=True, with_uv=True):
    t = (1 + 5 ** 0.5) / 2
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t),
         (0, 1, -t), (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4),
         (11, 10, 2), (10, 7, 6), (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9),
         (4, 9, 5), (2, 4, 11), (6, 2, 10), (8, 6, 7), (9, 8, 1)]
    v = [tuple(normalize(p)) for p in v]
    for _ in range(subdiv):
        cache, nf = {}, []

        def mid(a, b):
            key = (min(a, b), max(a, b))
            if key not in cache:
                cache[key] = len(v)
                v.append(tuple(normalize(np.add(v[a], v[b]) / 2)))
            return cache[key]

        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    n = np.asarray(v, dtype=np.float64)
    out = {"positions": f32(n * radius), "triangles": np.asarray(f, np.int32)}
    if with_normals:
        out["normals"] = f32(n)
    if with_uv:
        out["texcoords"] = f32(np.stack([np.arctan2(n[:, 0], n[:, 2]) / (2 * np.pi) + 0.5,
                                         np.arccos(np.clip(n[:, 1], -1, 1)) / np.pi], axis=1))
    return out


def cubesphere(steps, radius, rng):
    """Sphere made of quads (a cube blown up), with vertex colours."""
    pos, quads = [], []
    axes = [((1, 0, 0), (0, 1, 0), (0, 0, 1)), ((0, 1, 0), (0, 0, 1), (1, 0, 0)),
            ((0, 0, 1), (1, 0, 0), (0, 1, 0))]
    for u, v, w in axes:
        for sign in (1, -1):
            base = len(pos)
            for j in range(steps + 1):
                for i in range(steps + 1):
                    a, b = 2 * i / steps - 1, 2 * j / steps - 1
                    p = np.multiply(u, a) + np.multiply(v, b * sign) + np.multiply(w, sign)
                    pos.append(normalize(p) * radius)
            for j in range(steps):
                for i in range(steps):
                    k = base + j * (steps + 1) + i
                    quads.append((k, k + 1, k + steps + 2, k + steps + 1))
    pos = np.asarray(pos)
    colors = np.concatenate([rng.uniform(0.3, 1.0, size=(len(pos), 3)), np.ones((len(pos), 1))], 1)
    return {"positions": f32(pos), "normals": f32(pos / radius), "colors": f32(colors),
            "texcoords": f32(pos[:, :2] / radius * 0.5 + 0.5), "quads": np.asarray(quads, np.int32)}


def add_shapes(arrays, shapes):
    arrays["shapes.count"] = np.asarray([len(shapes)], np.int32)
    for i, s in enumerate(shapes):
        for k, v in s.items():
            arrays[f"shape{i}.{k}"] = v
        arrays[f"shape{i}.unsupported"] = np.zeros(2, np.int32)


def add_textures(arrays, textures):
    arrays["textures.count"] = np.asarray([len(textures)], np.int32)
    for i, (pixels, linear) in enumerate(textures):
        h, w = pixels.shape[:2]
        arrays[f"tex{i}.meta"] = np.asarray([w, h, int(linear)], np.int32)
        arrays[f"tex{i}." + ("pixelsf" if pixels.dtype == np.float32 else "pixelsb")] = pixels


def build_mixed():
    rng = np.random.default_rng(20260311)
    a = {"meta": np.asarray([1, 0, 0, 0], np.int32)}
    a["cameras"] = f32(np.stack([camera((0.0, 1.0, 3.4), (0, 0.25, 0)),
                                 camera((1.8, 1.2, 2.8), (0, 0.2, 0), aperture=0.02),
                                 camera((0.0, 1.5, 3.0), (0, 0.3, 0), lens=0.012, orthographic=True),
                                 camera((-1.5, 0.8, 2.6), (0, 0.3, 0), aspect=0.75, aperture=0.01)]))
    # textures: 0 colour (byte sRGB), 1 normal map (byte linear), 2 environment (float),
    # 3 roughness (byte sRGB), 4 emission (float)
    checker = (np.indices((16, 16)).sum(0) % 2)[..., None]
    col = (checker * rng.integers(120, 255, (16, 16, 3)) + (1 - checker) * rng.integers(20, 110, (16, 16, 3)))
    tex_color = np.concatenate([col, np.full((16, 16, 1), 255)], -1).astype(np.uint8)
    nrm = rng.normal(0, 0.25, (16, 16, 3))
    nrm[..., 2] = 1
    nrm /= np.linalg.norm(nrm, axis=-1, keepdims=True)
    tex_normal = np.concatenate([(nrm * 0.5 + 0.5) * 255, np.full((16, 16, 1), 255)], -1).astype(np.uint8)
    yy, xx = np.mgrid[0:16, 0:32]
    sky = np.stack([0.4 + 0.3 * np.cos(xx / 32 * 2 * np.pi), 0.5 + 0.2 * np.sin(yy / 16 * np.pi),
                    0.9 - 0.5 * yy / 16, np.ones_like(xx, dtype=np.float64)], -1)
    sky[3:5, 8:11, :3] = 30.0  # a small sun
    tex_env = f32(sky)
    tex_rough = np.concatenate([rng.integers(30, 200, (8, 8, 3)), np.full((8, 8, 1), 255)], -1).astype(np.uint8)
    tex_emit = f32(np.concatenate([rng.uniform(0.5, 1.0, (4, 4, 3)), np.ones((4, 4, 1))], -1))
    add_textures(a, [(tex_color, False), (tex_normal, True), (tex_env, True), (tex_rough, False),
                     (tex_emit, True)])

    tri_as_quad = quad_shape(0.3)
    tri_as_quad["quads"] = np.asarray([[0, 1, 2, 2], [0, 2, 3, 3]], np.int32)
    shapes = [quad_shape(2.0, uvscale=4.0),              # 0 floor
              icosphere(2, 0.25),                         # 1 triangles, normals + uv
              cubesphere(4, 0.25, rng),                   # 2 quads + vertex colours
              icosphere(1, 0.2, with_normals=False, with_uv=False),  # 3 facetted
              quad_shape(0.3),                            # 4 small quad
              tri_as_quad]                                # 5 degenerate quads
    add_shapes(a, shapes)

    mats = [material(MATTE, color=(0.9, 0.9, 0.9), color_tex=0, normal_tex=1),
            material(GLOSSY, color=(0.7, 0.2, 0.2), roughness=0.2),
            material(REFLECTIVE, color=(0.8, 0.8, 0.6), roughness=0.15),
            material(REFRACTIVE, color=(1, 1, 1), roughness=0.0, ior=1.5),
            material(TRANSPARENT, color=(0.6, 0.9, 0.7), roughness=0.1),
            material(VOLUMETRIC, color=(0.5, 0.5, 0.5), scattering=(0.7, 0.5, 0.3), scanisotropy=0.3,
                     trdepth=0.05),
            material(SUBSURFACE, color=(0.9, 0.7, 0.6), scattering=(0.8, 0.4, 0.3), roughness=0.2,
                     trdepth=0.02),
            material(GLTFPBR, color=(0.9, 0.8, 0.3), roughness=0.5, metallic=0.7, roughness_tex=3),
            material(REFLECTIVE, color=(0.9, 0.9, 0.9), roughness=0.0),
            material(MATTE, emission=(12, 11, 9), emission_tex=4),
            material(MATTE, color=(0.8, 0.8, 0.9), opacity=0.5),
            material(REFRACTIVE, color=(0.9, 1.0, 0.9), roughness=0.2, ior=1.33),
            material(TRANSPARENT, color=(0.9, 0.9, 1.0), roughness=0.0),
            material(MATTE, color=(0.2, 0.2, 0.2), emission=(5, 6, 8))]   # 13: triangle-mesh emitter
    a["materials.f"] = f32(np.stack([m[0] for m in mats]))
    a["materials.i"] = np.stack([m[1] for m in mats]).astype(np.int32)

    inst = [(frame(), 0, 0),
            (frame((-0.9, 0.25, 0.1)), 1, 1),
            (frame((-0.3, 0.25, -0.4)), 2, 2),
            (frame((0.3, 0.3, 0.2), scale=(1.0, 1.2, 0.8), rot_y=0.4), 1, 3),   # non-rigid
            (frame((0.9, 0.25, -0.1)), 1, 4),
            (frame((-0.6, 0.25, 0.7)), 2, 5),
            (frame((0.0, 0.2, 0.8)), 3, 6),
            (frame((0.6, 0.25, 0.75)), 1, 7),
            (frame((0.0, 0.6, -1.2), rot_x=np.pi / 2, scale=(3, 1, 2)), 4, 8),  # mirror wall
            (frame((0.0, 1.6, 0.3), rot_x=np.pi, scale=(1.5, 1, 1.5)), 4, 9),   # area light
            (frame((-0.45, 0.2, 1.25), rot_x=np.pi / 2, scale=(0.6, 1, 0.6)), 4, 10),  # half-opaque card
            (frame((1.3, 0.2, 0.5)), 3, 11),
            (frame((-1.3, 0.31, 0.5), rot_x=np.pi / 2), 5, 12),
            # an emissive TRIANGLE mesh (facetted icosphere, non-rigid frame): sample_triangle
            # in sample_lights and the triangle areas of make_lights
            (frame((0.95, 0.95, 0.55), scale=(0.45, 0.3, 0.45), rot_y=0.3), 3, 13)]
    a["instances.frame"] = f32(np.stack([i[0] for i in inst]))
    a["instances.ids"] = np.asarray([[i[1], i[2]] for i in inst], np.int32)
    # two environments (eval_environment sums them, yocto_scene.cpp:645-651); the second has
    # no texture: uniform-sphere light sampling and the 1 / 4 pi pdf (yocto_pathtrace.cpp:343, :414)
    a["environments.f"] = f32([np.concatenate([frame(rot_y=0.7), f32([0.8, 0.8, 0.8])]),
                               np.concatenate([frame(rot_y=-0.3, rot_x=0.2), f32([0.12, 0.1, 0.08])])])
    a["environments.i"] = np.asarray([[2], [-1]], np.int32)
    a["volumes.count"] = np.asarray([0], np.int32)
    a["vol_instances.f"] = np.zeros((0, 13), np.float32)
    a["vol_instances.i"] = np.zeros((0, 2), np.int32)
    a["sdfs.f"] = np.zeros((0, 19), np.float32)
    a["sdfs.i"] = np.zeros((0, 2), np.int32)
    return a


def build_sdf_small():
    a = {"meta": np.asarray([1, 0, 0, 0], np.int32)}
    a["cameras"] = f32(np.stack([camera((0.9, 0.8, 2.0), (0.1, 0.1, 0), aspect=1.5)]))
    yy, xx = np.mgrid[0:8, 0:16]
    sky = np.stack([0.5 + 0.2 * np.cos(xx / 16 * 2 * np.pi), 0.6 + 0 * xx, 0.9 - 0.4 * yy / 8,
                    np.ones_like(xx, dtype=np.float64)], -1)
    add_textures(a, [(f32(sky), True)])
    add_shapes(a, [])
    a["instances.frame"] = np.zeros((0, 12), np.float32)
    a["instances.ids"] = np.zeros((0, 2), np.int32)
    mats = [material(MATTE, color=(0.7, 0.7, 0.7)), material(MATTE, emission=(15, 15, 15)),
            material(GLOSSY, color=(0.4, 0.2, 0.1), roughness=0.1),
            material(REFLECTIVE, color=(0.9, 0.9, 0.9), roughness=0.3),
            material(MATTE, color=(0.2, 0.37, 0.5)), material(REFLECTIVE, color=(0.8, 0.8, 0.3), roughness=0.0),
            material(MATTE, color=(0.6, 0.3, 0.5), opacity=0.6)]
    a["materials.f"] = f32(np.stack([m[0] for m in mats]))
    a["materials.i"] = np.stack([m[1] for m in mats]).astype(np.int32)
    a["environments.f"] = f32([np.concatenate([frame(), f32([0.5, 0.5, 0.5])])])
    a["environments.i"] = np.asarray([[0]], np.int32)
    # a 16^3 grid holding the distance to a sphere of radius 0.3 centred in a 0.8 box
    n, res = 16, 0.05
    g = (np.arange(n) + 0.0) * res
    zz, yy, xx = np.meshgrid(g, g, g, indexing="ij")
    c = (n - 1) * res / 2
    a["volumes.count"] = np.asarray([1], np.int32)
    a["vol0.whd"] = np.asarray([n, n, n], np.int32)
    a["vol0.res"] = f32([res])
    a["vol0.data"] = f32(np.sqrt((xx - c) ** 2 + (yy - c) ** 2 + (zz - c) ** 2) - 0.3)
    a["vol_instances.f"] = f32([np.concatenate([frame((0.5, 0.1, -0.3)), f32([1.0])])])
    a["vol_instances.i"] = np.asarray([[0, 2]], np.int32)

    def sdf(type_, mat, o, params=(0, 0, 0, 0), whd=(0, 0, 0), **kw):
        return np.concatenate([frame(o, **kw), f32(params), f32(whd)]), (type_, mat)

    def sdf_at(type_, mat, world_center, whd, **kw):
        """world -> local is the FORWARD frame in the reference (yocto_sdfs.cpp:20): pick the
        frame origin so that the box [0, whd] sits around `world_center`."""
        m = frame((0, 0, 0), **kw).astype(np.float64)[:9].reshape(3, 3).T  # columns x, y, z
        o = np.asarray(whd, np.float64) / 2 - m @ np.asarray(world_center, np.float64)
        return sdf(type_, mat, tuple(o), whd=whd, **kw)

    sdfs = [sdf(SDF_PLANE, 0, (0, 0.25, 0)),                       # world -> local uses the forward frame
            sdf(SDF_SPHERE, 1, (0.3, -1.4, -0.2), params=(0.25, 0, 0, 0)),
            sdf(SDF_TORUS, 3, (0.6, 0.05, -0.5), params=(0.25, 0.08, 0, 0)),
            sdf(SDF_BOX, 4, (0.2, 0.1, 0.3), whd=(0.3, 0.3, 0.3)),
            sdf(SDF_BBOX, 5, (-0.3, -0.1, 0.5), params=(0.02, 0.2, 0.15, 0.2)),
            sdf(SDF_CAPPED_CONE, 6, (-0.2, -0.05, -0.4), params=(0.2, 0.2, 0.05, 0)),
            sdf(SDF_SPHERE, 2, (-0.9, -0.1, 0.2), params=(0.2, 0, 0, 0)),
            # an emissive box light with a rotated frame: the sdf branch of sample_lights
            # (inverse(frame), yocto_pathtrace.cpp:327-332) and of sample_lights_pdf (:382-394)
            sdf_at(SDF_BOX, 1, (-0.45, 0.75, 0.1), whd=(0.35, 0.25, 0.02), rot_y=0.5, rot_x=0.35)]
    a["sdfs.f"] = f32(np.stack([s[0] for s in sdfs]))
    a["sdfs.i"] = np.asarray([s[1] for s in sdfs], np.int32)
    return a


def build_dark():
    """No emitter at all (no emissive material, no environment): the light list is empty.
    The reference indexes an empty vector there (UB); the device ends the path (lights.cuh)."""
    a = {"meta": np.asarray([1, 0, 0, 0], np.int32)}
    a["cameras"] = f32(np.stack([camera((0.0, 1.0, 3.0), (0, 0.25, 0))]))
    add_textures(a, [])
    add_shapes(a, [quad_shape(2.0), icosphere(1, 0.4)])
    mats = [material(MATTE, color=(0.8, 0.8, 0.8)), material(GLOSSY, color=(0.6, 0.3, 0.2), roughness=0.3)]
    a["materials.f"] = f32(np.stack([m[0] for m in mats]))
    a["materials.i"] = np.stack([m[1] for m in mats]).astype(np.int32)
    inst = [(frame(), 0, 0), (frame((0, 0.4, 0)), 1, 1)]
    a["instances.frame"] = f32(np.stack([i[0] for i in inst]))
    a["instances.ids"] = np.asarray([[i[1], i[2]] for i in inst], np.int32)
    a["environments.f"] = np.zeros((0, 15), np.float32)
    a["environments.i"] = np.zeros((0, 1), np.int32)
    a["volumes.count"] = np.asarray([0], np.int32)
    a["vol_instances.f"] = np.zeros((0, 13), np.float32)
    a["vol_instances.i"] = np.zeros((0, 2), np.int32)
    a["sdfs.f"] = np.zeros((0, 19), np.float32)
    a["sdfs.i"] = np.zeros((0, 2), np.int32)
    return a


def frame_matrix(m, o):
    m = np.asarray(m, dtype=np.float64)
    return f32(np.concatenate([m[:, 0], m[:, 1], m[:, 2], np.asarray(o, dtype=np.float64)]))


def build_skewed():
    """Instances with strongly anisotropic and sheared frames (condition numbers 1 ... ~700): the
    object-space rays of the reference lose eps * cond of precision there, which is what the
    uploader's culling margin has to cover (csrc/upload.cuh cull_record)."""
    a = {"meta": np.asarray([1, 0, 0, 0], np.int32)}
    a["cameras"] = f32(np.stack([camera((0.0, 1.0, 3.4), (0, 0.25, 0))]))
    add_textures(a, [])
    add_shapes(a, [quad_shape(2.5), icosphere(2, 0.5), cubesphere(4, 0.35, np.random.default_rng(5))])
    mats = [material(MATTE, color=(0.7, 0.7, 0.7)), material(GLOSSY, color=(0.7, 0.3, 0.2), roughness=0.2),
            material(MATTE, color=(0.2, 0.6, 0.3)), material(REFLECTIVE, color=(0.8, 0.8, 0.9), roughness=0.05),
            material(MATTE, color=(0.9, 0.9, 0.9), emission=(6, 6, 5))]
    a["materials.f"] = f32(np.stack([m[0] for m in mats]))
    a["materials.i"] = np.stack([m[1] for m in mats]).astype(np.int32)
    ry = lambda t: np.array([[np.cos(t), 0, np.sin(t)], [0, 1, 0], [-np.sin(t), 0, np.cos(t)]])
    shear = np.array([[1.0, 0.9, 0.3], [0.0, 1.0, 0.6], [0.0, 0.0, 1.0]])
    inst = [(frame(), 0, 0),
            (frame_matrix(ry(0.3) @ np.diag([2.0, 0.02, 1.0]), (-0.8, 0.3, 0.2)), 1, 1),      # 100 : 1 pancake
            (frame_matrix(ry(-0.5) @ np.diag([0.015, 1.5, 0.015]), (0.1, 0.8, -0.3)), 1, 2),  # 100 : 1 needle
            (frame_matrix(ry(0.8) @ shear @ np.diag([1.0, 0.05, 0.6]), (0.9, 0.35, 0.3)), 2, 3),  # sheared slab
            (frame_matrix(shear.T @ np.diag([0.01, 1.0, 1.0]), (-0.2, 0.6, 0.9)), 2, 1),      # sheared sheet
            (frame_matrix(ry(0.2) @ np.diag([3.0, 0.004, 3.0]), (0.0, 1.9, 0.0)), 1, 4),      # 750 : 1 light
            (frame((0.5, 0.25, 0.9)), 2, 2)]
    a["instances.frame"] = f32(np.stack([i[0] for i in inst]))
    a["instances.ids"] = np.asarray([[i[1], i[2]] for i in inst], np.int32)
    a["environments.f"] = f32([np.concatenate([frame(), f32([0.3, 0.3, 0.35])])])
    a["environments.i"] = np.asarray([[-1]], np.int32)
    a["volumes.count"] = np.asarray([0], np.int32)
    a["vol_instances.f"] = np.zeros((0, 13), np.float32)
    a["vol_instances.i"] = np.zeros((0, 2), np.int32)
    a["sdfs.f"] = np.zeros((0, 19), np.float32)
    a["sdfs.i"] = np.zeros((0, 2), np.int32)
    return a


def main():
    import __graft_entry__ as entry
    import util

    vpt = entry.load_package()
    vpt.archive.save(os.path.join(HERE, "mixed.vptscene"), build_mixed())
    vpt.archive.save(os.path.join(HERE, "sdf_small.vptscene"), build_sdf_small())
    vpt.archive.save(os.path.join(HERE, "dark.vptscene"), build_dark())
    vpt.archive.save(os.path.join(HERE, "skewed.vptscene"), build_skewed())
    print("wrote mixed.vptscene, sdf_small.vptscene, dark.vptscene, skewed.vptscene")
    if "--scenes" in sys.argv:
        return
    if not util.have_ref():
        sys.exit("oracle/_ref/ref_render is not built (needs /root/reference): goldens unchanged")
    for scene, shader, spp, res, bounces in GOLDEN_RENDERS:
        out = util.golden_name(scene, shader, spp, res, bounces)
        subprocess.run([util.REF_RENDER, os.path.join(HERE, scene + ".vptscene"), out, "--shader",
                        shader, "--samples", str(spp), "--resolution", str(res), "--bounces",
                        str(bounces), "--dump-state"], check=True, stdout=subprocess.DEVNULL)
        print("golden", os.path.basename(out))
    # first hits of a fixed ray batch, and the host structures (bvh, lights, initial rngs)
    rays = util.camera_rays(0, 0, np.random.default_rng(7), 4096)
    for scene in ("mixed", "sdf_small"):
        rp = os.path.join(HERE, "_rays.vptarch")
        vpt.archive.save(rp, {"rays": rays})
        subprocess.run([util.REF_RENDER, os.path.join(HERE, scene + ".vptscene"),
                        os.path.join(HERE, f"{scene}_hits.vptarch"), "--rays", rp, "--stmaxiter",
                        "450"], check=True, stdout=subprocess.DEVNULL)
        os.remove(rp)
    subprocess.run([util.REF_RENDER, os.path.join(HERE, "mixed.vptscene"),
                    os.path.join(HERE, "mixed_host.vptarch"), "--dump-host", "--resolution", "64"],
                   check=True, stdout=subprocess.DEVNULL)
    print("golden mixed_hits, sdf_small_hits, mixed_host")


if __name__ == "__main__":
    main()
