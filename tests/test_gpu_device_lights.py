"""GPU: make_lights on the device (yocto_pathtrace.cpp:983-1049).

vpt_cuda_upload_scene_device_lights enumerates the lights in the reference's order and builds
every elements_cdf in HBM (areas / texel weights in parallel, running sums in the reference's
sequential float order). Gates: the light list and every CDF equal the reference's make_lights
bit for bit (its dump in tests/golden/mixed_host.vptarch, and the host build the CPU suite pins to
that dump, on every other scene), and renders that use the device-built lights stay bit-exact."""
import os

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu

SCENES = ["mixed", "sdf_small", "03_volume", "01_surface", "05_head1ss", "06_gridsdf", "07_sdfunction"]


def same_bits(a, b):
    return a.shape == b.shape and not (a.view(np.uint32) != b.view(np.uint32)).any()


def test_device_lights_equal_reference_dump(vpt):
    ref = vpt.archive.load(os.path.join(util.GOLDEN, "mixed_host.vptarch"))
    sc = vpt.binding.Scene(util.scene_path("mixed"))
    st = sc.use_device_lights()
    lights = sc.device_lights()
    assert st["lights"] == len(lights) == ref["lights.ids"].shape[0]
    for i, l in enumerate(lights):
        assert [l["instance"], l["environment"]] == list(ref["lights.ids"][i][:2])
        assert same_bits(l["cdf"], ref[f"lights.cdf{i}"]), i
    sc.close()


@pytest.mark.parametrize("scene", SCENES)
def test_device_lights_equal_host_make_lights(vpt, scene):
    path = util.scene_path(scene)
    if path is None:
        pytest.skip(f"{scene}.vptscene not built")
    sc = vpt.binding.Scene(path)
    host = sc.lights()
    st = sc.use_device_lights()
    dev = sc.device_lights()
    assert len(dev) == len(host) == st["lights"]
    for i, (h, d) in enumerate(zip(host, dev)):
        assert (h["instance"], h["environment"], h["sdf"]) == (d["instance"], d["environment"], d["sdf"])
        assert same_bits(h["cdf"], d["cdf"]), (scene, i, int((h["cdf"] != d["cdf"]).argmax()))
    assert st["cdf_elements"] == sum(l["cdf"].size for l in host if l["sdf"] < 0)
    sc.close()


@pytest.mark.parametrize("scene,shader,spp,res,bounces", [
    ("mixed", "volpathtrace", 2, 64, 8), ("mixed", "pathtrace", 2, 64, 4),
    ("sdf_small", "implicit", 2, 64, 4)])
def test_render_with_device_lights_is_bit_exact(vpt, scene, shader, spp, res, bounces):
    ref = util.oracle_render(vpt, scene, shader, spp, res, bounces=bounces)
    assert ref is not None
    sc = vpt.binding.Scene(util.scene_path(scene))
    sc.use_device_lights()
    r = vpt.binding.Render(sc, vpt.binding.make_params(shader=shader, samples=spp, resolution=res,
                                                      bounces=bounces))
    r.trace(spp)
    assert not util.bit_mismatch(r.image(normalized=False), ref["image_sum"]).any()
    assert np.array_equal(r.rngs(), ref["rngs_after"])
    r.close()
    sc.close()
