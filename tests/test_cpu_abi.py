"""CPU suite: the C-ABI boundary and the failure behaviour without a device.

  * every function include/vpt_cuda.h and include/vpt_host.h declare is exported by the
    built library (no compute calls here);
  * with no CUDA device the product path fails loudly (VPT_ERR_NO_DEVICE): there is no CPU
    fallback, and nothing under volumetric-path-tracer_b200/ refers to oracle/;
  * the fdlibm restatements the device uses for atan / atan2 / acos equal this box's libm.
"""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

import util

PKG = os.path.join(util.ROOT, "volumetric-path-tracer_b200")


def declared(header, prefix):
    text = open(os.path.join(util.ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    return sorted(set(re.findall(r"\b(" + prefix + r"\w+)\s*\(", text)))


@pytest.mark.parametrize("header,prefix,lib", [("vpt_cuda.h", "vpt_cuda_", "libvpt_cuda.so"),
                                               ("vpt_host.h", "vpth_", "libvpt_host.so")])
def test_library_exports_every_declared_symbol(vpt, header, prefix, lib):
    vpt.binding.libs()  # loads libvpt_cuda.so first (libvpt_host.so links against it)
    names = declared(header, prefix)
    assert len(names) >= 15, names
    handle = ctypes.CDLL(os.path.join(PKG, lib))
    missing = [n for n in names if not hasattr(handle, n)]
    assert not missing, f"{lib} does not export {missing}"


def test_north_star_entry_points_present(vpt):
    names = declared("vpt_cuda.h", "vpt_cuda_")
    for n in ("vpt_cuda_upload_scene", "vpt_cuda_trace_samples", "vpt_cuda_download_image"):
        assert n in names


def test_status_codes_and_struct_sizes(vpt):
    """vpt_bvh_node is layout-compatible with yocto::bvh_node (32 bytes: bbox3f, int32 start,
    int16 num, int8 axis, bool internal); the ctypes mirror of vpt_params matches the header."""
    b = vpt.binding
    assert b.BVH_NODE_DTYPE.itemsize == 32
    header = open(os.path.join(util.ROOT, "include", "vpt_cuda.h")).read()
    for name, value in (("VPT_OK", 0), ("VPT_ERR_INVALID", -1), ("VPT_ERR_UNSUPPORTED", -2),
                        ("VPT_ERR_NO_DEVICE", -3), ("VPT_ERR_CUDA", -4), ("VPT_ERR_OUT_OF_MEMORY", -5)):
        assert re.search(name + r"\s*=\s*" + str(value) + r"\b", header), name
    fields = re.search(r"typedef struct vpt_params \{(.*?)\} vpt_params;", header, re.S).group(1)
    fields = re.sub(r"/\*.*?\*/", "", fields, flags=re.S)
    fields = re.sub(r"//[^\n]*", "", fields)
    n_fields = len(re.findall(r";", fields))
    assert n_fields == len(b.Params._fields_)


def test_no_device_fails_loudly(vpt):
    """On a box without a GPU every compute entry point returns VPT_ERR_NO_DEVICE with a
    message; nothing is rendered on the CPU instead."""
    b = vpt.binding
    if b.device_count() > 0:
        pytest.skip("a CUDA device is present")
    sc = b.Scene(util.scene_path("mixed"))
    r = b.Render(sc, b.make_params(shader="color", samples=1, resolution=32))
    with pytest.raises(b.VptError) as e:
        r.trace(1)
    assert e.value.args and "no CUDA device" in str(e.value) and "-3" in str(e.value)
    with pytest.raises(b.VptError):
        sc.intersect(np.zeros((4, 8), np.float32))
    with pytest.raises(b.VptError):
        b.set_device(0)
    assert r.samples == 0
    r.close()
    sc.close()


def test_product_tree_does_not_reach_the_oracle():
    """Only tests/, smoke() and bench.py's baseline legs may touch oracle/."""
    hits = []
    for base, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".cu", ".cuh", ".cpp", ".h", ".py")) or f == "Makefile":
                text = open(os.path.join(base, f), errors="ignore").read()
                # comments may name the exporter that writes .vptscene files; code may not
                text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
                text = re.sub(r"(//|#)[^\n]*", "", text)
                if re.search(r"oracle/|ref_render|libyocto_ref|/root/reference", text):
                    hits.append(os.path.join(base, f))
    assert not hits, hits


def test_device_libm_restatements_match_glibc(vpt):
    """What the device runs instead of libm — vatan / vatan2 / vacos (csrc/vmath.cuh, fdlibm) and
    g_sinf / g_cosf / g_expf / g_logf / g_powf (csrc/glibc_libm.cuh, glibc's flt-32 algorithms
    with the FMA placement of its x86-64 FMA build) — against this box's libm: 0 mismatching bit
    patterns over every 257th float (16.7 M arguments per function, NaNs / infinities /
    subnormals / huge arguments included), pow additionally over 600 k random pairs, and
    sincosf == (sinf, cosf). The full sweep (stride 1, all 2^32 floats) also gives 0 and takes
    ~4 minutes: tests/hostsim/hostsim --selftest-math 2000000 1."""
    vpt.binding.libs()
    r = subprocess.run(["make", "-s", "-f", os.path.join(util.ROOT, "tests", "hostsim", "Makefile")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    r = subprocess.run([os.path.join(util.ROOT, "tests", "hostsim", "hostsim"), "--selftest-math",
                        "200000", "257"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "math selftest: 0 mismatches" in r.stdout
