"""GPU: the ypathtrace command line (reference flags, apps/ypathtrace/ypathtrace.cpp:307-341) end to end:
archive -> make_bvh / make_lights / make_state -> pathtrace_samples through the C ABI -> get_render ->
image file. The float image it writes equals the oracle's render divided by the sample count."""
import os
import subprocess

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu
CLI = os.path.join(util.ROOT, "volumetric-path-tracer_b200", "ypathtrace")


def read_pfm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split())
        assert float(f.readline()) < 0  # little endian
        data = np.frombuffer(f.read(), dtype="<f4").reshape(h, w, 3)
    return data[::-1]  # stored bottom row first


@pytest.mark.parametrize("scene,shader,bounces", [("mixed", "pathtrace", 4), ("sdf_small", "implicit", 4),
                                                  ("mixed", "eyelight", 4)])
def test_cli_output_equals_oracle(vpt, tmp_path, scene, shader, bounces):
    spp, res = 4, 96
    ref = util.oracle_render(vpt, scene, shader, spp, res, bounces=bounces)
    if ref is None:
        pytest.skip("oracle binary absent")
    out = str(tmp_path / "out.pfm")
    r = subprocess.run([CLI, "--scene", util.scene_path(scene), "--output", out, "--shader", shader,
                        "--samples", str(spp), "--resolution", str(res), "--bounces", str(bounces)],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    got = read_pfm(out)
    want = ref["image_sum"][..., :3] * np.float32(1.0 / spp)  # get_render: sum * (1 / samples)
    assert got.shape == want.shape
    assert not util.bit_mismatch(np.ascontiguousarray(got), np.ascontiguousarray(want)).any()


def test_cli_png_and_batching(vpt, tmp_path):
    """--batch changes how many samples go into one device call, never the result."""
    outs = []
    for batch in (1, 3, 64):
        out = str(tmp_path / f"b{batch}.pfm")
        r = subprocess.run([CLI, "--scene", util.scene_path("mixed"), "--output", out, "--shader",
                            "volpathtrace", "--samples", "6", "--resolution", "64", "--bounces", "8",
                            "--batch", str(batch)], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        outs.append(read_pfm(out))
    assert np.array_equal(outs[0], outs[1]) and np.array_equal(outs[0], outs[2])
    png = str(tmp_path / "o.png")
    r = subprocess.run([CLI, "--scene", util.scene_path("mixed"), "--output", png, "--shader", "color",
                        "--samples", "1", "--resolution", "64"], capture_output=True, text=True)
    assert r.returncode == 0 and open(png, "rb").read(8) == b"\x89PNG\r\n\x1a\n"


def test_cli_device_lights_changes_nothing(vpt, tmp_path):
    """--device-lights: make_lights runs on the GPU; the render is the same image bit for bit."""
    outs = []
    for extra in ([], ["--device-lights"]):
        out = str(tmp_path / f"l{len(extra)}.pfm")
        r = subprocess.run([CLI, "--scene", util.scene_path("mixed"), "--output", out, "--shader",
                            "volpathtrace", "--samples", "4", "--resolution", "64", "--bounces", "8",
                            *extra], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        assert ("build lights (device)" in r.stdout) == bool(extra)
        outs.append(read_pfm(out))
    assert np.array_equal(outs[0], outs[1])


def test_cli_rejects_what_the_reference_rejects(tmp_path):
    r = subprocess.run([CLI, "--scene", util.scene_path("mixed"), "--output", str(tmp_path / "x.pfm"),
                        "--shader", "nosuchshader"], capture_output=True, text=True)
    assert r.returncode != 0
    r = subprocess.run([CLI, "--scene", util.scene_path("mixed"), "--output", str(tmp_path / "x.pfm"),
                        "--interactive"], capture_output=True, text=True)
    assert r.returncode != 0 and "unavailable on this box" in (r.stdout + r.stderr)


@pytest.mark.parametrize("scene,shader", [("mixed", "pathtrace"), ("mixed", "color"), ("sdf_small", "implicit")])
def test_device_output_stage_equals_reference_bytes(vpt, tmp_path, scene, shader):
    """get_render + rgb_to_srgb + float_to_byte on the device. The bytes equal what
    the reference's save_image encodes for an 8-bit format (rgb_to_srgb(vector<vec4b>&, ...),
    yocto_image.cpp:870-874), through the binding and in the CLI's .ppm output."""
    spp, res = 4, 96
    ref = util.oracle_render(vpt, scene, shader, spp, res)
    if ref is None or "image_srgb8" not in ref:
        pytest.skip("oracle binary absent")
    sc = vpt.binding.Scene(util.scene_path(scene))
    r = vpt.binding.Render(sc, vpt.binding.make_params(shader=shader, samples=spp, resolution=res))
    r.trace(spp)
    got = r.image_srgb8()
    r.close()
    sc.close()
    assert got.shape == ref["image_srgb8"].shape
    assert np.array_equal(got, ref["image_srgb8"])
    assert got[..., :3].std() > 5, "image should not be flat"
    out = str(tmp_path / "o.ppm")
    c = subprocess.run([CLI, "--scene", util.scene_path(scene), "--output", out, "--shader", shader,
                        "--samples", str(spp), "--resolution", str(res)], capture_output=True, text=True)
    assert c.returncode == 0, c.stdout + c.stderr
    with open(out, "rb") as f:
        assert f.readline().strip() == b"P6"
        w, h = map(int, f.readline().split())
        assert f.readline().strip() == b"255"
        rgb = np.frombuffer(f.read(), dtype=np.uint8).reshape(h, w, 3)
    assert np.array_equal(rgb, ref["image_srgb8"][..., :3])
