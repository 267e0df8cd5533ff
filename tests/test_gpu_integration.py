"""GPU: the reference-side adapter (integration/yocto_pathtrace_cuda.cpp) — the reference's own
`pathtrace_samples` / `get_render` signatures, compiled against the reference's headers, fed by
the reference's own `load_scene` from scene JSON, `tesselate_surfaces`, `make_bvh`,
`make_lights`, `make_state` — against the same program linked with the reference's CPU tracer.

Binaries (oracle/Makefile, built where /root/reference exists, carried to the GPU box):
  ypathtrace_headless_{cpu,cuda}  headless run_offline + --dump of the float render / sums / rngs
  ypathtrace_ref_{cpu,cuda}       the reference's apps/ypathtrace/ypathtrace.cpp itself (GL cut)
"""
import os
import subprocess

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu

REF = os.path.join(util.ROOT, "oracle", "_ref")
JSON = os.path.join(util.ROOT, "scenes", "_json")
SCENES = {"03_volume": os.path.join(JSON, "03_volume", "volume.json"),
          "sdf_mini": os.path.join(JSON, "sdf_mini", "sdf_mini.json")}


def need(*paths):
    for p in paths:
        if not os.path.exists(p):
            pytest.skip(f"{os.path.relpath(p, util.ROOT)} not built (needs /root/reference at build time)")


def run(binary, *args):
    r = subprocess.run([os.path.join(REF, binary), *map(str, args)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return r


@pytest.mark.parametrize("scene,shader,spp,res,bounces", [
    ("03_volume", "volpathtrace", 3, 360, 16), ("03_volume", "pathtrace", 2, 240, 4),
    ("03_volume", "eyelight", 1, 240, 4), ("sdf_mini", "implicit", 3, 240, 4),
    ("sdf_mini", "implicit_normal", 1, 240, 4)])
def test_adapter_equals_reference_from_json(vpt, tmp_path, scene, shader, spp, res, bounces):
    """Same scene JSON, same loaders and host builders, same CLI flags: the CUDA adapter's
    get_render image, accumulation sums and per-pixel generators equal the reference CPU
    tracer's, bit for bit."""
    need(os.path.join(REF, "ypathtrace_headless_cpu"), os.path.join(REF, "ypathtrace_headless_cuda"),
         SCENES[scene])
    a, b = str(tmp_path / "cpu.vptarch"), str(tmp_path / "cuda.vptarch")
    args = ("--scene", SCENES[scene], "--shader", shader, "--samples", spp, "--resolution", res,
            "--bounces", bounces)
    run("ypathtrace_headless_cpu", *args, "--dump", a)
    run("ypathtrace_headless_cuda", *args, "--dump", b)
    cpu, gpu = vpt.archive.load(a), vpt.archive.load(b)
    assert list(cpu["meta"]) == list(gpu["meta"])
    assert cpu["render"][..., :3].max() > 0
    for key in ("render", "image_sum"):
        bad = util.bit_mismatch(cpu[key], gpu[key])
        assert not bad.any(), f"{key}: {int(bad.sum())} pixels differ"
    assert np.array_equal(cpu["rngs_after"], gpu["rngs_after"])


def test_reference_cli_source_runs_on_the_adapter(tmp_path):
    """The reference's own ypathtrace.cpp (viewer removed, nothing else) linked with the adapter
    writes the same PNG bytes as the same source linked with the reference's CPU tracer, with
    the command line of scripts/run.sh:3 at a reduced size."""
    need(os.path.join(REF, "ypathtrace_ref_cpu"), os.path.join(REF, "ypathtrace_ref_cuda"),
         SCENES["03_volume"])
    a, b = str(tmp_path / "cpu.png"), str(tmp_path / "cuda.png")
    args = ("--scene", SCENES["03_volume"], "--shader", "volpathtrace", "--samples", 4,
            "--resolution", 240, "--bounces", 64)
    run("ypathtrace_ref_cpu", *args, "--output", a)
    out = run("ypathtrace_ref_cuda", *args, "--output", b)
    assert "render image" in out.stdout  # the reference's progress bar, per sample
    assert open(a, "rb").read() == open(b, "rb").read()


def test_interactive_is_refused_headless():
    """--interactive needs the OpenGL viewer: unavailable on this box; print_fatal exits non-zero
    (yocto_cli.h:365-384 convention)."""
    need(os.path.join(REF, "ypathtrace_headless_cuda"), SCENES["sdf_mini"])
    r = subprocess.run([os.path.join(REF, "ypathtrace_headless_cuda"), "--scene", SCENES["sdf_mini"],
                        "--shader", "implicit_normal", "--samples", "1", "--resolution", "64",
                        "--interactive"], capture_output=True, text=True)
    assert r.returncode != 0 and "unavailable on this box" in (r.stdout + r.stderr)
