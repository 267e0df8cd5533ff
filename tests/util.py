"""Shared helpers of the test-suite: oracle invocation and image metrics.

The oracle is oracle/_ref/ref_render — the UNMODIFIED reference CPU path built
from /root/reference (it travels to the GPU box as a binary). Where it is
absent the tests fall back to the committed golden outputs under tests/golden/.
"""
from __future__ import annotations

import os
import subprocess
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_RENDER = os.path.join(ROOT, "oracle", "_ref", "ref_render")
GOLDEN = os.path.join(ROOT, "tests", "golden")
BUILT = os.path.join(ROOT, "scenes", "_built")


def have_ref() -> bool:
    return os.path.exists(REF_RENDER)


def scene_path(name: str):
    for base in (BUILT, GOLDEN):
        p = os.path.join(base, name + ".vptscene")
        if os.path.exists(p):
            return p
    return None


def golden_name(scene, shader, samples, resolution, bounces, extra=()):
    tag = "_".join([scene, shader, f"{resolution}px", f"{samples}spp", f"b{bounces}", *extra])
    return os.path.join(GOLDEN, tag + ".vptarch")


def oracle_render(vpt, scene, shader, samples, resolution, bounces=4, extra_args=(),
                  allow_golden=True):
    """Reference render -> dict with image_sum [h,w,4] (+ rngs_after). Live oracle if the binary
    is here, else the committed golden for exactly these settings, else None."""
    gold = golden_name(scene, shader, samples, resolution, bounces,
                       tuple(a.strip("-") for a in extra_args))
    if have_ref():
        with tempfile.TemporaryDirectory() as tmp:
            out = os.path.join(tmp, "ref.vptarch")
            cmd = [REF_RENDER, scene_path(scene), out, "--shader", shader, "--samples",
                   str(samples), "--resolution", str(resolution), "--bounces", str(bounces),
                   "--dump-state", *extra_args]
            subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL)
            return vpt.archive.load(out)
    if allow_golden and os.path.exists(gold):
        return vpt.archive.load(gold)
    return None


def oracle_hits(vpt, scene, rays, instance=-1, maxiter=450):
    if not have_ref():
        return None
    with tempfile.TemporaryDirectory() as tmp:
        rp, out = os.path.join(tmp, "rays.vptarch"), os.path.join(tmp, "hits.vptarch")
        vpt.archive.save(rp, {"rays": np.ascontiguousarray(rays, dtype=np.float32)})
        cmd = [REF_RENDER, scene_path(scene), out, "--rays", rp, "--stmaxiter", str(maxiter)]
        if instance >= 0:
            cmd += ["--instance", str(instance)]
        subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL)
        return vpt.archive.load(out)


def oracle_host_dump(vpt, scene, resolution=720):
    if not have_ref():
        return None
    with tempfile.TemporaryDirectory() as tmp:
        out = os.path.join(tmp, "host.vptarch")
        subprocess.run([REF_RENDER, scene_path(scene), out, "--dump-host", "--resolution",
                        str(resolution)], check=True, stdout=subprocess.DEVNULL)
        return vpt.archive.load(out)


def bit_mismatch(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Per-pixel mask of pixels whose float bits differ in any channel."""
    return (np.ascontiguousarray(a).view(np.uint32) != np.ascontiguousarray(b).view(np.uint32)
            ).any(axis=-1)


def luminance(img):
    return 0.2126 * img[..., 0] + 0.7152 * img[..., 1] + 0.0722 * img[..., 2]


def box_blur3(img):
    pad = np.pad(img, ((1, 1), (1, 1), (0, 0)), mode="edge")
    h, w = img.shape[:2]
    return sum(pad[i:i + h, j:j + w] for i in range(3) for j in range(3)) / 9.0


def tonemap_srgb(linear):
    x = np.clip(linear, 0, 1)
    return np.where(x <= 0.0031308, 12.92 * x, 1.055 * np.power(x, 1 / 2.4) - 0.055)


def rmse(a, b):
    return float(np.sqrt(np.mean((a - b) ** 2)))


def camera_rays(width, height, rng, n):
    """Seeded rays fanning out of a point in front of the test scenes (y up, looking down -z)."""
    o = np.tile(np.array([[0.0, 0.3, 1.2]], dtype=np.float32), (n, 1))
    o += rng.uniform(-0.4, 0.4, size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d[:, 2] = -np.abs(d[:, 2]) - 0.5
    d[:, 1] -= 0.3
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.concatenate([o, d, np.full((n, 1), 1e-4, np.float32),
                           np.full((n, 1), np.finfo(np.float32).max, np.float32)], axis=1)
    return rays.astype(np.float32)
