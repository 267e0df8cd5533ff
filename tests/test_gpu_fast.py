"""GPU: the statistical build (volumetric-path-tracer_b200/fast, -DVPT_FAST: FMA contraction,
approximate division / square root, CUDA libm, 1-FFMA slab tests). north_star's tolerance for
stochastic shaders is what it must meet - NOT bit-exactness, which the default build keeps:
  * equal-spp renders vs the reference: mean luminance within 1 %, RMSE <= 0.02 on 3x3-blurred sRGB;
  * the reference's own check image of tests/03_volume within the same tolerance;
  * first hits equal the reference's except for a recorded, tiny share of grazing rays.
bench.py reports its throughput next to the exact build's (`fast` key): the price of bit-exactness."""
import os
import subprocess
import sys

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu
FAST_DIR = os.path.join(util.ROOT, "volumetric-path-tracer_b200", "fast")
WORKER = os.path.join(util.ROOT, "tests", "fast_worker.py")


def run_fast(*args):
    if not os.path.exists(os.path.join(FAST_DIR, "libvpt_cuda.so")):
        pytest.skip("fast build absent")
    env = dict(os.environ, VPT_LIB_DIR=FAST_DIR)
    r = subprocess.run([sys.executable, WORKER, *map(str, args)], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]


@pytest.mark.parametrize("scene,shader,bounces", [("03_volume", "volpathtrace", 64), ("01_surface", "pathtrace", 4),
                                                  ("mixed", "volpathtrace", 16), ("mixed", "naive", 8),
                                                  ("07_sdfunction", "implicit", 4), ("sdf_small", "implicit", 4)])
def test_fast_build_matches_oracle_statistically(vpt, tmp_path, scene, shader, bounces):
    spp, res = 64, 160
    ref = util.oracle_render(vpt, scene, shader, spp, res, bounces)
    if ref is None:
        pytest.skip("no oracle output for this case")
    out = str(tmp_path / "img.npy")
    run_fast("render", scene, shader, spp, res, bounces, out)
    img = np.load(out)[..., :3] / spp
    want = ref["image_sum"][..., :3] / spp
    lum_a, lum_b = util.luminance(np.clip(img, 0, 4)).mean(), util.luminance(np.clip(want, 0, 4)).mean()
    assert abs(lum_a / lum_b - 1) <= 0.01, f"mean luminance {lum_a} vs oracle {lum_b}"
    err = util.rmse(util.box_blur3(util.tonemap_srgb(img)), util.box_blur3(util.tonemap_srgb(want)))
    assert err <= 0.02, f"rmse {err}"


def test_fast_build_check_image_03_volume(vpt, tmp_path):
    gold = os.path.join(util.ROOT, "tests", "golden", "check_03_volume_720_256.npy")
    want = np.load(gold).astype(np.float32) / 255.0
    out = str(tmp_path / "img.npy")
    run_fast("render", "03_volume", "volpathtrace", 256, 720, 64, out)
    img = util.tonemap_srgb(np.load(out)[..., :3] / 256)
    assert img.shape == want.shape
    assert abs(util.luminance(img).mean() / util.luminance(want).mean() - 1) <= 0.01
    assert util.rmse(util.box_blur3(img), util.box_blur3(want)) <= 0.02


@pytest.mark.parametrize("scene", ["03_volume", "mixed"])
def test_fast_build_first_hits_except_recorded_ties(vpt, tmp_path, scene):
    """Deterministic gate of north_star 'except for recorded ties': with approximate arithmetic a
    grazing ray may pick the neighbouring primitive or miss; the share is recorded and bounded,
    every other ray agrees in {hit, instance, element} and to 1e-4 relative in distance."""
    rays = util.camera_rays(0, 0, np.random.default_rng(7), 50000)
    ref = util.oracle_hits(vpt, scene, rays)
    if ref is None:
        pytest.skip("oracle binary absent")
    rp, out = str(tmp_path / "rays.npy"), str(tmp_path / "hits.npz")
    np.save(rp, rays)
    run_fast("hits", scene, rp, out)
    got = np.load(out)
    differ = (got["ids"] != ref["hits.ids"]).any(axis=1)
    share = differ.mean()
    print(f"{scene}: {int(differ.sum())} of {len(rays)} rays pick another primitive ({100 * share:.4f} %)")
    assert share <= 2e-3
    same = ~differ & (ref["hits.ids"][:, 0] == 1)
    d0, d1 = ref["hits.f"][same, 2], got["uvd"][same, 2]
    assert np.abs(d1 - d0).max() <= 1e-4 * np.abs(d0).max()
