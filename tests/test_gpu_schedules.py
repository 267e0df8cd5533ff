"""GPU: the alternative SIMT schedules of the extend and light-pdf stages give the same image bit for bit.

VPT_WF_POOL=1 selects k_wf_extend_pool (csrc/wavefront.cuh): a warp owns a pool of 96 rays in shared
memory and runs one kind of traversal operation at a time for the rays that hold it. Per ray the
operations and their order are those of the shipped kernel, so accumulation images and generator
states must be identical. It is measured slower than the shipped kernel on B200
and therefore off by default; this test keeps it honest.
"""
import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu


def render(vpt, scene, shader, res, bounces, spp):
    b = vpt.binding
    sc = b.Scene(scene)
    r = b.Render(sc, b.make_params(shader=shader, samples=64, resolution=res, bounces=bounces))
    r.trace(spp)
    img, rng = r.image(normalized=False).copy(), r.rngs().copy()
    r.close(), sc.close()
    return img, rng


@pytest.mark.parametrize("scene,shader,res,bounces", [("mixed", "pathtrace", 160, 8),
                                                      ("mixed", "volpathtrace", 96, 16),
                                                      ("03_volume", "volpathtrace", 320, 64)])
@pytest.mark.parametrize("limit", [None, "3"])
def test_pooled_extend_is_bit_identical(vpt, monkeypatch, scene, shader, res, bounces, limit):
    path = util.scene_path(scene)
    if path is None:
        pytest.skip(f"{scene} not built")
    monkeypatch.delenv("VPT_WF_POOL", raising=False)
    want, want_rng = render(vpt, path, shader, res, bounces, 5)
    monkeypatch.setenv("VPT_WF_POOL", "1")
    if limit:  # three stack entries in shared memory: nearly every walk uses the spill column
        monkeypatch.setenv("VPT_WF_POOL_LIMIT", limit)
    got, got_rng = render(vpt, path, shader, res, bounces, 5)
    assert not util.bit_mismatch(got, want).any()
    assert np.array_equal(got_rng, want_rng)
    assert np.abs(want).sum() > 0


@pytest.mark.parametrize("scene,shader,res,bounces", [("mixed", "pathtrace", 160, 8), ("mixed", "volpathtrace", 96, 16),
                                                      ("sdf_small", "implicit", 128, 4),
                                                      ("07_sdfunction", "implicit", 320, 4)])
def test_persistent_light_pdf_is_bit_identical(vpt, monkeypatch, scene, shader, res, bounces):
    """VPT_WF_PDF_PERSISTENT=1: sample_lights_pdf as a per-lane state machine with lane refill
    (k_wf_lightpdf_persistent). Measured slower than one thread per queued ray, hence opt-in."""
    path = util.scene_path(scene)
    if path is None:
        pytest.skip(f"{scene} not built")
    monkeypatch.delenv("VPT_WF_PDF_PERSISTENT", raising=False)
    want, want_rng = render(vpt, path, shader, res, bounces, 5)
    monkeypatch.setenv("VPT_WF_PDF_PERSISTENT", "1")
    got, got_rng = render(vpt, path, shader, res, bounces, 5)
    assert not util.bit_mismatch(got, want).any()
    assert np.array_equal(got_rng, want_rng)
    assert np.abs(want).sum() > 0


@pytest.mark.parametrize("scene,shader,res,bounces", [("mixed", "volpathtrace", 128, 16), ("01_surface", "pathtrace", 256, 4)])
def test_extend_register_builds_are_bit_identical(vpt, monkeypatch, scene, shader, res, bounces):
    """VPT_WF_ROOMY=0 / 1: k_wf_extend_persistent compiled for 6 blocks per SM (80 registers) and for 5
    (96 registers); the host picks by the scene's triangle count, the image cannot depend on it."""
    path = util.scene_path(scene)
    if path is None:
        pytest.skip(f"{scene} not built")
    monkeypatch.setenv("VPT_WF_ROOMY", "0")
    want, want_rng = render(vpt, path, shader, res, bounces, 5)
    monkeypatch.setenv("VPT_WF_ROOMY", "1")
    got, got_rng = render(vpt, path, shader, res, bounces, 5)
    assert not util.bit_mismatch(got, want).any()
    assert np.array_equal(got_rng, want_rng)
