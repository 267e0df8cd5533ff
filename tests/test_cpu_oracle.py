"""CPU suite (-m "not gpu"): the oracle against the committed golden vectors, the host side
(BVH build, lights, per-pixel generators, archive I/O) against the reference's own
structures, and the device headers compiled for the host (tests/hostsim) against the
goldens. Nothing here launches a kernel.

Golden vectors (tests/golden/*.vptarch) were written by oracle/_ref/ref_render, the
UNMODIFIED reference CPU path, with tests/golden/make_fixtures.py.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import util

sys.path.insert(0, os.path.join(util.ROOT, "tests", "golden"))
import make_fixtures  # noqa: E402

HOSTSIM = os.path.join(util.ROOT, "tests", "hostsim", "hostsim")


def golden(vpt, scene, shader, spp, res, bounces):
    path = util.golden_name(scene, shader, spp, res, bounces)
    assert os.path.exists(path), f"golden missing: {path}"
    return vpt.archive.load(path)


@pytest.fixture(scope="session")
def hostsim(vpt):
    """Host build of csrc/*.cuh (the same headers the kernels are made of)."""
    vpt.binding.libs()
    r = subprocess.run(["make", "-s", "-f", os.path.join(util.ROOT, "tests", "hostsim", "Makefile")],
                       capture_output=True, text=True)
    if r.returncode != 0:
        pytest.fail("hostsim did not build:\n" + r.stderr[-2000:])
    return HOSTSIM


def run_tool(binary, scene, out, *args):
    subprocess.run([binary, util.scene_path(scene), out, *map(str, args)], check=True,
                   stdout=subprocess.DEVNULL)


# ---- the oracle is pinned to the committed goldens ------------------------------------------------
@pytest.mark.parametrize("case", make_fixtures.GOLDEN_RENDERS, ids=lambda c: "-".join(map(str, c)))
def test_oracle_reproduces_golden(vpt, tmp_path, case):
    """ref_render built from /root/reference today == the committed golden, bit for bit."""
    if not util.have_ref():
        pytest.skip("oracle/_ref/ref_render not built (needs /root/reference)")
    scene, shader, spp, res, bounces = case
    out = str(tmp_path / "o.vptarch")
    run_tool(util.REF_RENDER, scene, out, "--shader", shader, "--samples", spp, "--resolution", res,
             "--bounces", bounces, "--dump-state")
    got, ref = vpt.archive.load(out), golden(vpt, *case)
    assert not util.bit_mismatch(got["image_sum"], ref["image_sum"]).any()
    assert np.array_equal(got["rngs_after"], ref["rngs_after"])


def test_oracle_is_deterministic_across_thread_counts(vpt, tmp_path):
    """parallel_for vs --noparallel give the same image: pixels own their streams (C14)."""
    if not util.have_ref():
        pytest.skip("oracle/_ref/ref_render not built")
    a, b = str(tmp_path / "a.vptarch"), str(tmp_path / "b.vptarch")
    args = ("--shader", "pathtrace", "--samples", 2, "--resolution", 48)
    run_tool(util.REF_RENDER, "mixed", a, *args)
    run_tool(util.REF_RENDER, "mixed", b, *args, "--noparallel")
    assert not util.bit_mismatch(vpt.archive.load(a)["image_sum"], vpt.archive.load(b)["image_sum"]).any()


# ---- device headers on the host == goldens ---------------------------------------------------------
@pytest.mark.parametrize("case", make_fixtures.GOLDEN_RENDERS, ids=lambda c: "-".join(map(str, c)))
def test_device_code_on_host_matches_golden(vpt, hostsim, tmp_path, case):
    """All shaders, every pixel and every generator state: the arithmetic the kernels are
    compiled from replays the reference exactly (the GPU differs only through libm, see
    test_gpu_parity.py)."""
    scene, shader, spp, res, bounces = case
    out = str(tmp_path / "h.vptarch")
    run_tool(hostsim, scene, out, "--shader", shader, "--samples", spp, "--resolution", res,
             "--bounces", bounces)
    got, ref = vpt.archive.load(out), golden(vpt, *case)
    bad = util.bit_mismatch(got["image_sum"], ref["image_sum"])
    assert not bad.any(), f"{int(bad.sum())} pixels differ"
    assert np.array_equal(got["rngs_after"], ref["rngs_after"])


@pytest.mark.parametrize("camera", [1, 2, 3])
def test_cameras_on_host_match_oracle(vpt, hostsim, tmp_path, camera):
    """Thin-lens, orthographic and portrait cameras: device headers on the host vs the live oracle."""
    if not util.have_ref():
        pytest.skip("oracle/_ref/ref_render not built")
    a, b = str(tmp_path / "a.vptarch"), str(tmp_path / "b.vptarch")
    args = ("--shader", "pathtrace", "--samples", 2, "--resolution", 64, "--camera", camera)
    run_tool(util.REF_RENDER, "mixed", a, *args, "--dump-state")
    run_tool(hostsim, "mixed", b, *args)
    ref, got = vpt.archive.load(a), vpt.archive.load(b)
    assert ref["image_sum"].shape == got["image_sum"].shape
    assert not util.bit_mismatch(got["image_sum"], ref["image_sum"]).any()
    assert np.array_equal(got["rngs_after"], ref["rngs_after"])


def test_first_hits_on_skewed_frames_match_oracle(vpt, hostsim, tmp_path):
    """Grazing and ordinary rays against 100 : 1 scaled and sheared instances: device headers on the
    host vs the live oracle, ids and float bits."""
    if not util.have_ref():
        pytest.skip("oracle/_ref/ref_render not built")
    rays = util.camera_rays(0, 0, np.random.default_rng(17), 20000)
    rp, out = str(tmp_path / "rays.vptarch"), str(tmp_path / "hits.vptarch")
    vpt.archive.save(rp, {"rays": rays})
    run_tool(hostsim, "skewed", out, "--rays", rp)
    got, ref = vpt.archive.load(out), util.oracle_hits(vpt, "skewed", rays)
    assert ref["hits.ids"][:, 0].sum() > 1000
    assert np.array_equal(got["hits.ids"], ref["hits.ids"])
    assert not util.bit_mismatch(got["hits.f"], ref["hits.f"]).any()


@pytest.mark.parametrize("scene", ["mixed", "sdf_small"])
def test_first_hits_on_host_match_golden(vpt, hostsim, tmp_path, scene):
    rays = util.camera_rays(0, 0, np.random.default_rng(7), 4096)
    rp, out = str(tmp_path / "rays.vptarch"), str(tmp_path / "hits.vptarch")
    vpt.archive.save(rp, {"rays": rays})
    run_tool(hostsim, scene, out, "--rays", rp, "--stmaxiter", 450)
    got, ref = vpt.archive.load(out), vpt.archive.load(os.path.join(util.GOLDEN, f"{scene}_hits.vptarch"))
    for key in ref:
        if ref[key].dtype == np.float32:
            assert not util.bit_mismatch(got[key], ref[key]).any(), key
        else:
            assert np.array_equal(got[key], ref[key]), key
    if scene == "mixed":
        assert ref["hits.ids"][:, 0].sum() > 1000


@pytest.mark.parametrize("scene,shader,spp,res,bounces,least", [("mixed", "volpathtrace", 2, 64, 8, 1000),
                                                                ("skewed", "pathtrace", 2, 64, 4, 1000)])
def test_instance_culling_never_hides_a_hit(vpt, hostsim, tmp_path, scene, shader, spp, res, bounces, least):
    """bvh.cuh skips an instance whose enlarged world box the ray misses (the reference transforms
    the ray and tests the root box in object space). `--cull-check` walks every culled instance the
    way the reference would: none of them may hold a hit. The render must also stay the golden.
    `skewed` holds frames with condition numbers up to ~750 (100 : 1 scales, shear): the margin
    scales with the condition number (upload.cuh cull_record)."""
    import re
    out = str(tmp_path / "h.vptarch")
    r = subprocess.run([hostsim, util.scene_path(scene), out, "--shader", shader, "--samples", str(spp),
                        "--resolution", str(res), "--bounces", str(bounces), "--cull-check"],
                       check=True, capture_output=True, text=True)
    m = re.search(r"cull check: (\d+) instances culled, (\d+) of them", r.stdout)
    assert m, r.stdout
    assert int(m.group(1)) > least and int(m.group(2)) == 0, r.stdout
    got, ref = vpt.archive.load(out), golden(vpt, scene, shader, spp, res, bounces)
    assert not util.bit_mismatch(got["image_sum"], ref["image_sum"]).any()
    assert np.array_equal(got["rngs_after"], ref["rngs_after"])


@pytest.mark.parametrize("mode", [1, 2])
def test_sdf_field_culling_is_exact(vpt, hostsim, tmp_path, mode):
    """sdf.cuh eval_sdf_scene_culled (a distance field whose conservative lower bound is above
    the running minimum is not evaluated; opt-in, the shipped kernels scan all fields): every
    skipped field is evaluated anyway and must lie strictly above the running minimum, and the
    image equals the reference's golden bit for bit."""
    import re
    scene, shader, spp, res, bounces = "sdf_small", "implicit", 2, 64, 4
    out = str(tmp_path / "h.vptarch")
    r = subprocess.run([hostsim, util.scene_path(scene), out, "--shader", shader, "--samples", str(spp),
                        "--resolution", str(res), "--bounces", str(bounces), "--sdf-cull", str(mode),
                        "--sdf-cull-check"], check=True, capture_output=True, text=True)
    m = re.search(r"sdf cull check: (\d+) fields skipped, (\d+) of them", r.stdout)
    assert m, r.stdout
    assert int(m.group(1)) > 1000 and int(m.group(2)) == 0, r.stdout
    got, ref = vpt.archive.load(out), golden(vpt, scene, shader, spp, res, bounces)
    assert not util.bit_mismatch(got["image_sum"], ref["image_sum"]).any()
    assert np.array_equal(got["rngs_after"], ref["rngs_after"])


# ---- host structures -----------------------------------------------------------------------------
def test_host_bvh_is_the_reference_bvh(vpt):
    """make_bvh (host/vpt_bvh.cpp) vs yocto's make_bvh: nodes byte for byte, primitive order equal."""
    ref = vpt.archive.load(os.path.join(util.GOLDEN, "mixed_host.vptarch"))
    sc = vpt.binding.Scene(util.scene_path("mixed"))
    counts = sc.counts()
    for shape in [-1, *range(counts["shapes"])]:
        name = "scene" if shape < 0 else f"shape{shape}"
        nodes, prims = sc.bvh(shape)
        assert nodes.tobytes() == ref[f"bvh.{name}.nodes"].tobytes(), name
        assert np.array_equal(prims, ref[f"bvh.{name}.prims"]), name
    sc.close()


def test_lights_and_cdfs_match_reference(vpt):
    ref = vpt.archive.load(os.path.join(util.GOLDEN, "mixed_host.vptarch"))
    sc = vpt.binding.Scene(util.scene_path("mixed"))
    lights = sc.lights()
    # quad area light, TRIANGLE-mesh emitter (80 triangle areas), textured environment (texel
    # weights), texture-less environment (empty CDF)
    assert len(lights) == ref["lights.ids"].shape[0] == 4
    assert [len(l["cdf"]) for l in lights] == [1, 80, 32 * 16, 0]
    for i, l in enumerate(lights):
        assert [l["instance"], l["environment"], l["sdf"]][:2] == list(ref["lights.ids"][i][:2])
        assert not (l["cdf"].view(np.uint32) != ref[f"lights.cdf{i}"].view(np.uint32)).any()
    sc.close()


def test_make_state_generators(vpt):
    """Per-pixel PCG32 seeding (yocto_pathtrace.cpp:1116-1121): one meta generator with seed
    1301081, each pixel seeded with a draw from it. Known answer for pixel 0 plus the dump."""
    ref = vpt.archive.load(os.path.join(util.GOLDEN, "mixed_host.vptarch"))
    sc = vpt.binding.Scene(util.scene_path("mixed"))
    r = vpt.binding.Render(sc, vpt.binding.make_params(shader="color", samples=1, resolution=64))
    assert [r.width, r.height] == list(ref["state.size"])
    rngs = r.rngs()
    assert np.array_equal(rngs, ref["state.rngs"])
    assert (int(rngs[0, 0]), int(rngs[0, 1])) == (10473825460370060855, 1450249601)
    r.close()
    sc.close()


def test_archive_round_trip(vpt, tmp_path):
    arrays = {"f": np.arange(12, dtype=np.float32).reshape(3, 4), "i": np.asarray([-1, 2], np.int32),
              "b": np.arange(6, dtype=np.uint8).reshape(1, 2, 3), "q": np.asarray([2 ** 63 + 5], np.uint64),
              "empty": np.zeros((0, 13), np.float32)}
    p = str(tmp_path / "a.vptarch")
    vpt.archive.save(p, arrays)
    back = vpt.archive.load(p)
    assert list(back) == list(arrays)
    for k in arrays:
        assert back[k].dtype == arrays[k].dtype and back[k].shape == arrays[k].shape
        assert np.array_equal(back[k], arrays[k])
    with open(p, "r+b") as f:
        f.write(b"NOTVPT")
    with pytest.raises(ValueError):
        vpt.archive.load(p)


def test_fixture_scenes_rebuild_identically(vpt, tmp_path):
    """The committed fixtures are exactly what make_fixtures.py writes (seeded numpy)."""
    for name, build in (("mixed", make_fixtures.build_mixed), ("sdf_small", make_fixtures.build_sdf_small)):
        p = str(tmp_path / (name + ".vptscene"))
        vpt.archive.save(p, build())
        with open(p, "rb") as a, open(os.path.join(util.GOLDEN, name + ".vptscene"), "rb") as b:
            assert a.read() == b.read(), name


def test_malformed_scene_is_rejected(vpt, tmp_path):
    a = make_fixtures.build_mixed()
    del a["materials.i"]
    p = str(tmp_path / "bad.vptscene")
    vpt.archive.save(p, a)
    with pytest.raises(vpt.binding.VptError):
        vpt.binding.Scene(p)
    with pytest.raises(vpt.binding.VptError):
        vpt.binding.Scene(str(tmp_path / "missing.vptscene"))


def test_oracle_reproduces_reference_check_image(vpt, tmp_path):
    """The reference ships its own render of tests/03_volume (check/lowres/03_volume_720_256.jpg,
    made by scripts/run.sh: volpathtrace, 256 spp, 720 px, 64 bounces). The oracle at the same
    settings must land on it: mean luminance within 1 %, RMSE <= 0.02 on 3x3-blurred sRGB
    (measured: ratio 0.9998, RMSE 0.0092; the residue is JPEG q75 + sampling noise)."""
    gold = os.path.join(util.GOLDEN, "check_03_volume_720_256.npy")
    if not util.have_ref() or util.scene_path("03_volume") is None:
        pytest.skip("needs oracle/_ref/ref_render and scenes/_built/03_volume.vptscene")
    out = str(tmp_path / "o.vptarch")
    run_tool(util.REF_RENDER, "03_volume", out, "--shader", "volpathtrace", "--samples", 256,
             "--resolution", 720, "--bounces", 64)
    img = util.tonemap_srgb(vpt.archive.load(out)["image_sum"][..., :3] / 256)
    want = np.load(gold).astype(np.float32) / 255.0
    assert img.shape == want.shape
    assert abs(util.luminance(img).mean() / util.luminance(want).mean() - 1) <= 0.01
    assert util.rmse(util.box_blur3(img), util.box_blur3(want)) <= 0.02
