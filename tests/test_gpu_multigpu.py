"""GPU: multi-GPU rendering THROUGH THE PRODUCT — pathtrace_params.gpus in the
C++ facade, `ypathtrace --gpus N`, peers + the P2P reduce inside vpt_cuda_download_image — and the
output stage's display tonemap.

On a one-GPU box the ranks share the device (VPT_GPUS_OVERSUBSCRIBE=1): host threads, replicas,
generator offsets, the sample split and the reduce kernel are all the same code; with two or more
GPUs visible the same tests also run across devices (P2P loads over NVLink).
"""
import os
import subprocess

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu
CLI = os.path.join(util.ROOT, "volumetric-path-tracer_b200", "ypathtrace")


@pytest.fixture()
def oversubscribe(vpt, monkeypatch):
    if vpt.binding.device_count() < 2:
        monkeypatch.setenv("VPT_GPUS_OVERSUBSCRIBE", "1")
    return vpt.binding.device_count()


@pytest.mark.parametrize("scene,shader,bounces", [("mixed", "volpathtrace", 8), ("sdf_small", "implicit", 4)])
def test_two_gpu_render_is_the_sum_of_its_ranks(vpt, oversubscribe, scene, shader, bounces):
    """gpus = 2, 6 + 5 samples in two calls: the downloaded image equals, bit for bit, rank 0's
    image (6 samples of the reference streams: 3 + 3) plus rank 1's (5 samples of the streams
    1 << 40 draws ahead: 3 + 2), each rendered alone on one device; the split is sample_share."""
    b, sh = vpt.binding, vpt.sharding
    sc = b.Scene(util.scene_path(scene))
    params = b.make_params(shader=shader, samples=64, resolution=96, bounces=bounces)
    multi = b.Render(sc, params, gpus=2)
    multi.trace(6)   # 3 + 3
    multi.trace(5)   # 3 + 2
    assert multi.samples == 11
    got = multi.image(normalized=False)
    got8 = multi.image_srgb8()
    parts = []
    for rank, n in ((0, 6), (1, 5)):
        r = b.Render(sc, params)
        r.reset(sh.rank_rng_offset(rank))
        r.trace(n)
        parts.append(r.image(normalized=False))
        r.close()
    want = parts[0] + parts[1]
    assert not util.bit_mismatch(got, want).any()
    # the reference's streams are rank 0's: its generators come back from the root
    assert np.abs(got).sum() > 0
    # the 8-bit output divides the SUM by the total sample count
    single = np.clip((util.tonemap_srgb(want[..., :3] / 11) * 256).astype(np.int64), 0, 255)
    assert np.abs(got8[..., :3].astype(np.int64) - single).max() <= 1
    multi.reset(0)
    multi.trace(2)  # 1 + 1 after a reset: replicas restart too
    assert multi.samples == 2
    a = b.Render(sc, params)
    a.trace(1)
    c = b.Render(sc, params)
    c.reset(sh.rank_rng_offset(1))
    c.trace(1)
    assert not util.bit_mismatch(multi.image(False), a.image(False) + c.image(False)).any()
    for r in (multi, a, c):
        r.close()
    sc.close()


def test_more_gpus_than_visible_is_an_error(vpt, monkeypatch):
    monkeypatch.delenv("VPT_GPUS_OVERSUBSCRIBE", raising=False)
    b = vpt.binding
    sc = b.Scene(util.scene_path("mixed"))
    r = b.Render(sc, b.make_params(shader="pathtrace", samples=4, resolution=32), gpus=b.device_count() + 1)
    with pytest.raises(Exception, match="GPUs requested"):
        r.trace(2)
    r.close()
    sc.close()


def test_cli_gpus_flag(vpt, oversubscribe, tmp_path):
    """`ypathtrace --gpus 2`: same flags otherwise; the float image is the two ranks' sum / spp.
    With deterministic 4-draw shaders the split image still differs from one GPU's (rank 1 jitters
    with other draws), so the check is against the two single-device renders."""
    b, sh = vpt.binding, vpt.sharding
    out = str(tmp_path / "o.pfm")
    spp, res = 8, 96
    env = dict(os.environ)
    c = subprocess.run([CLI, "--scene", util.scene_path("mixed"), "--output", out, "--shader", "pathtrace",
                        "--samples", str(spp), "--resolution", str(res), "--gpus", "2", "--batch", "8"],
                       capture_output=True, text=True, env=env)
    assert c.returncode == 0, c.stdout + c.stderr
    assert "2 GPUs" in c.stdout
    with open(out, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split())
        f.readline()
        got = np.frombuffer(f.read(), dtype="<f4").reshape(h, w, 3)[::-1]
    sc = b.Scene(util.scene_path("mixed"))
    params = b.make_params(shader="pathtrace", samples=64, resolution=res)
    parts = []
    for rank in (0, 1):
        r = b.Render(sc, params)
        r.reset(sh.rank_rng_offset(rank))
        r.trace(4)
        parts.append(r.image(normalized=False))
        r.close()
    sc.close()
    want = (parts[0] + parts[1])[..., :3] * np.float32(1.0 / spp)
    assert not util.bit_mismatch(np.ascontiguousarray(got), np.ascontiguousarray(want)).any()


def test_reduce_images_entry_point(vpt):
    """vpt_cuda_reduce_images (dst += sum of peer buffers, the kernel the downloads use), called
    on the device images of two states through the C ABI."""
    import ctypes as C

    b = vpt.binding
    cuda, _ = b.libs()
    sc = b.Scene(util.scene_path("mixed"))
    params = b.make_params(shader="pathtrace", samples=8, resolution=64)
    r0, r1 = b.Render(sc, params), b.Render(sc, params)
    r1.reset(1 << 40)
    r0.trace(2), r1.trace(3)
    a, c = r0.image(False), r1.image(False)
    peers = (C.c_void_p * 1)(r1.device_image_ptr())
    cuda.vpt_cuda_reduce_images.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, C.c_size_t]
    assert cuda.vpt_cuda_reduce_images(r0.device_image_ptr(), peers, 1, a.size) == 0
    torch = pytest.importorskip("torch")
    buf = torch.empty(a.size, dtype=torch.float32, device="cuda")
    r0.export_image(buf.data_ptr())  # the device image itself (the host mirror is cached)
    torch.cuda.synchronize()
    assert not util.bit_mismatch(buf.cpu().numpy().reshape(a.shape), a + c).any()
    assert cuda.vpt_cuda_reduce_images(r0.device_image_ptr(), peers, 1, a.size + 1) != 0  # not a multiple of 4
    r0.close(), r1.close(), sc.close()


@pytest.mark.parametrize("key,exposure,filmic", [("image_tonemapped8_filmic_e1.5", 1.5, True),
                                                 ("image_tonemapped8_e-0.75", -0.75, False)])
def test_display_tonemap_equals_reference_bytes(vpt, key, exposure, filmic):
    """tonemap(hdr, exposure, filmic) (yocto_color.h:306-316) + float_to_byte on the
    device, as the reference's viewer converts every frame (ypathtrace.cpp:170, :186)."""
    spp, res = 4, 96
    ref = util.oracle_render(vpt, "mixed", "pathtrace", spp, res)
    if ref is None or key not in ref:
        pytest.skip("oracle binary absent")
    sc = vpt.binding.Scene(util.scene_path("mixed"))
    r = vpt.binding.Render(sc, vpt.binding.make_params(shader="pathtrace", samples=spp, resolution=res))
    r.trace(spp)
    got = r.image_tonemapped8(exposure, filmic)
    assert np.array_equal(got, ref[key])
    assert not np.array_equal(got, r.image_srgb8())
    r.close(), sc.close()


def test_cli_display_loop(vpt, tmp_path):
    """--display / --display-every: the viewer's loop without a window. The preview is the
    one-sample render at resolution / pratio blown up; the last frame equals the final image
    behind the tonemap."""
    disp, out = str(tmp_path / "d.ppm"), str(tmp_path / "o.ppm")
    c = subprocess.run([CLI, "--scene", util.scene_path("mixed"), "--output", out, "--shader", "pathtrace",
                        "--samples", "8", "--resolution", "96", "--batch", "2", "--display", disp,
                        "--display-every", "4", "--pratio", "4", "--exposure", "0.5", "--filmic"],
                       capture_output=True, text=True)
    assert c.returncode == 0, c.stdout + c.stderr
    assert "preview: 24x16, 1 sample" in c.stdout

    def ppm(path):
        with open(path, "rb") as f:
            assert f.readline().strip() == b"P6"
            w, h = map(int, f.readline().split())
            f.readline()
            return np.frombuffer(f.read(), dtype=np.uint8).reshape(h, w, 3)

    sc = vpt.binding.Scene(util.scene_path("mixed"))
    r = vpt.binding.Render(sc, vpt.binding.make_params(shader="pathtrace", samples=8, resolution=96))
    r.trace(8)
    assert np.array_equal(ppm(disp), r.image_tonemapped8(0.5, True)[..., :3])
    assert np.array_equal(ppm(out), r.image_srgb8()[..., :3])
    r.close(), sc.close()
