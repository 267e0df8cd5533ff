"""GPU: edge cases of the boundary — flags, unsupported inputs, degenerate scenes."""
import os
import sys

import numpy as np
import pytest

import util

sys.path.insert(0, os.path.join(util.ROOT, "tests", "golden"))
import make_fixtures  # noqa: E402

pytestmark = pytest.mark.gpu


def render(vpt, scene_path, **kw):
    sc = vpt.binding.Scene(scene_path)
    spp = kw.get("samples", 2)
    r = vpt.binding.Render(sc, vpt.binding.make_params(**kw))
    r.trace(spp)
    img, rng = r.image(normalized=False), r.rngs()
    r.close()
    sc.close()
    return img, rng


@pytest.mark.parametrize("pipeline", ["mega", "wavefront"])
def test_noimplicit_mis_flag_bit_exact(vpt, pipeline):
    """shade_implicit with pure BSDF sampling (yocto_pathtrace.cpp:425-535, params.noimplicit_mis)."""
    ref = util.oracle_render(vpt, "sdf_small", "implicit", 3, 128, extra_args=("--noimplicitmis",))
    if ref is None:
        pytest.skip("oracle binary absent")
    img, rng = render(vpt, util.scene_path("sdf_small"), shader="implicit", samples=3, resolution=128,
                      noimplicit_mis=True, pipeline=pipeline)
    assert not util.bit_mismatch(img, ref["image_sum"]).any()
    assert np.array_equal(rng, ref["rngs_after"])


def test_spheretrace_iteration_cap_bit_exact(vpt):
    """--stmaxiter: rays that run out of iterations miss, exactly as in the reference."""
    ref = util.oracle_render(vpt, "sdf_small", "implicit_normal", 1, 128, extra_args=("--stmaxiter", "12"))
    if ref is None:
        pytest.skip("oracle binary absent")
    img, _ = render(vpt, util.scene_path("sdf_small"), shader="implicit_normal", samples=1,
                    resolution=128, spheretrace_maxiter=12)
    assert not util.bit_mismatch(img, ref["image_sum"]).any()


def test_one_bounce_and_many_bounces_bit_exact(vpt):
    for bounces in (1, 33):
        ref = util.oracle_render(vpt, "mixed", "volpathtrace", 2, 96, bounces=bounces)
        if ref is None:
            pytest.skip("oracle binary absent")
        img, rng = render(vpt, util.scene_path("mixed"), shader="volpathtrace", samples=2,
                          resolution=96, bounces=bounces)
        assert not util.bit_mismatch(img, ref["image_sum"]).any(), bounces
        assert np.array_equal(rng, ref["rngs_after"])


@pytest.mark.parametrize("shader", ["pathtrace", "volpathtrace", "naive"])
@pytest.mark.parametrize("pipeline", ["mega", "wavefront"])
def test_scene_without_emitters_renders_black(vpt, shader, pipeline):
    """No emissive material and no environment: the light list is empty. The reference indexes
    an empty vector there (undefined behaviour on the host); the device must not read out of
    bounds: light sampling yields a zero direction, which ends the path (lights.cuh), so the
    lit shaders render black with alpha = 1 on geometry, and the context stays usable."""
    path = util.scene_path("dark")
    img, _ = render(vpt, path, shader=shader, samples=4, resolution=96, pipeline=pipeline)
    assert np.isfinite(img).all()
    assert (img[..., :3] == 0).all()
    assert img[..., 3].max() == 4 and img[..., 3].min() == 0
    # the device is still healthy: a normal render right after
    img2, _ = render(vpt, util.scene_path("mixed"), shader="color", samples=1, resolution=64)
    assert img2[..., :3].max() > 0


def test_bounces_outside_the_packed_range_are_rejected(vpt):
    sc = vpt.binding.Scene(util.scene_path("mixed"))
    for bounces in (0, 65536):
        r = vpt.binding.Render(sc, vpt.binding.make_params(shader="pathtrace", samples=1,
                                                           resolution=32, bounces=bounces))
        with pytest.raises(Exception, match="bounces"):
            r.trace(1)
        r.close()
    sc.close()


def test_point_and_line_primitives_are_rejected(vpt, tmp_path):
    """The device path supports triangles and quads (all the reference's scenes); anything else
    fails at upload with VPT_ERR_UNSUPPORTED instead of rendering something different."""
    a = make_fixtures.build_mixed()
    a["shape3.unsupported"] = np.asarray([5, 0], np.int32)  # five point primitives
    p = str(tmp_path / "points.vptscene")
    vpt.archive.save(p, a)
    with pytest.raises(vpt.binding.VptError) as e:
        sc = vpt.binding.Scene(p)
        r = vpt.binding.Render(sc, vpt.binding.make_params(shader="color", samples=1, resolution=32))
        r.trace(1)
    assert "-2" in str(e.value) and "not supported" in str(e.value)


def test_scene_without_instances_renders_the_environment(vpt, tmp_path):
    """No shapes, no sdfs: every path is a miss; pathtrace returns the environment."""
    a = make_fixtures.build_mixed()
    for k in list(a):
        if k.startswith("shape"):
            del a[k]
    a["shapes.count"] = np.asarray([0], np.int32)
    a["instances.frame"] = np.zeros((0, 12), np.float32)
    a["instances.ids"] = np.zeros((0, 2), np.int32)
    p = str(tmp_path / "empty.vptscene")
    vpt.archive.save(p, a)
    img, _ = render(vpt, p, shader="pathtrace", samples=2, resolution=48)
    assert np.isfinite(img).all() and img[..., :3].mean() > 0.1
    if util.have_ref():
        import subprocess
        out = str(tmp_path / "ref.vptarch")
        subprocess.run([util.REF_RENDER, p, out, "--shader", "pathtrace", "--samples", "2",
                        "--resolution", "48"], check=True, stdout=subprocess.DEVNULL)
        assert not util.bit_mismatch(img, vpt.archive.load(out)["image_sum"]).any()


def test_state_reset_restarts_the_streams(vpt):
    sc = vpt.binding.Scene(util.scene_path("mixed"))
    r = vpt.binding.Render(sc, vpt.binding.make_params(shader="pathtrace", samples=4, resolution=64))
    r.trace(4)
    first = r.image(normalized=False).copy()
    r.reset()
    assert r.samples == 0
    r.trace(2)
    r.trace(2)
    assert not util.bit_mismatch(first, r.image(normalized=False)).any()
    r.close()
    sc.close()


@pytest.mark.parametrize("env", [{"VPT_WF_PIPES": "1"}, {"VPT_WF_PIPES": "3"}, {"VPT_WF_PIPES": "4"},
                                 {"VPT_WF_DRAIN": "0"}, {"VPT_WF_NO_PERSISTENT": "1"}],
                         ids=lambda e: "-".join(f"{k}={v}" for k, v in e.items()))
def test_wavefront_scheduling_never_changes_the_image(vpt, monkeypatch, env):
    """Pipes, drain hand-over and the persistent extend only reschedule work: whatever the setting,
    the image sums and generator states stay the reference's (yocto_pathtrace.cpp:1052-1092)."""
    ref = util.oracle_render(vpt, "mixed", "volpathtrace", 2, 64, bounces=8)
    assert ref is not None
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    img, rng = render(vpt, util.scene_path("mixed"), shader="volpathtrace", samples=2, resolution=64,
                      bounces=8, pipeline="wavefront")
    assert not util.bit_mismatch(img, ref["image_sum"]).any()
    assert np.array_equal(rng, ref["rngs_after"])


@pytest.mark.parametrize("suspend", ["0", "8", "32"])
def test_parked_rays_never_change_the_image(vpt, monkeypatch, suspend):
    """Sphere-tracing warps park their unfinished rays for the next launch (wavefront.cuh); the march
    itself is unchanged, so any threshold gives the reference's image (yocto_pathtrace.cpp:289-307)."""
    ref = util.oracle_render(vpt, "sdf_small", "implicit", 2, 64, bounces=4)
    assert ref is not None
    monkeypatch.setenv("VPT_WF_SUSPEND", suspend)
    img, rng = render(vpt, util.scene_path("sdf_small"), shader="implicit", samples=2, resolution=64,
                      bounces=4, pipeline="wavefront")
    assert not util.bit_mismatch(img, ref["image_sum"]).any()
    assert np.array_equal(rng, ref["rngs_after"])


def test_image_view_is_the_image_without_a_copy(vpt):
    """vpth_render_image_view: the host mirror itself (read-only numpy view), equal to image(False);
    a second render call refreshes it in place."""
    b = vpt.binding
    sc = b.Scene(util.scene_path("mixed"))
    r = b.Render(sc, b.make_params(shader="pathtrace", samples=16, resolution=64))
    r.trace(2)
    view = r.image_view()
    assert view.shape == (r.height, r.width, 4) and not view.flags.writeable
    assert not util.bit_mismatch(view, r.image(normalized=False)).any()
    before = view.copy()
    r.trace(2)
    again = r.image_view()
    assert again.ctypes.data == view.ctypes.data
    assert util.bit_mismatch(again, before).any()
    assert not util.bit_mismatch(again, r.image(normalized=False)).any()
    r.close(), sc.close()
