import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def vpt():
    """The vpt_b200 package (ctypes binding + archive tools)."""
    return entry.load_package()


@pytest.fixture(scope="session")
def root():
    return ROOT


def scene_path(name):
    """Built scene archive (scenes/_built, produced by build()) or committed fixture."""
    for base in (os.path.join(ROOT, "scenes", "_built"), os.path.join(ROOT, "tests", "golden")):
        p = os.path.join(base, name + ".vptscene")
        if os.path.exists(p):
            return p
    return None


REF_RENDER = os.path.join(ROOT, "oracle", "_ref", "ref_render")


def have_ref():
    return os.path.exists(REF_RENDER)


# ---- host libm variant (VERDICT r01 "parity gaps" (v)) ---------------------------------------------
# The bit-exact gates compare the GPU with the LIVE oracle, which calls this host's libm; the device
# restates glibc's FMA+AVX2 flt-32 variants (csrc/glibc_libm.cuh). On a host whose dynamic linker
# picks other variants (no FMA / AVX2, another glibc) the oracle itself rounds a few arguments
# differently, and a mismatch there says nothing about the kernels. tests/hostsim --selftest-math
# (device restatements compiled for the host vs this host's libm) decides: if it reports a
# mismatch, every test that holds the GPU to the live oracle's bits is skipped with that reason
# and the statistical gates (test_stochastic_shaders_match_oracle, test_check_image_03_volume,
# tests/test_gpu_fast.py) are what this host checks. GPU-vs-GPU identities and the committed
# goldens' CPU tests do not depend on it.
_EXACT_VS_LIVE_ORACLE = ("bit_exact", "bit_for_bit", "equals_reference", "equals_oracle", "reference_dump",
                         "equals_the_reference", "adapter_equals", "reference_cli_source")
_libm_state = {}


def host_libm_is_the_restated_variant():
    if "ok" not in _libm_state:
        import subprocess
        hostsim = os.path.join(ROOT, "tests", "hostsim", "hostsim")
        ok = True
        if os.path.exists(hostsim):
            try:  # every 65 537th float bit pattern per function + 20 000 random pairs: about a second
                r = subprocess.run([hostsim, "--selftest-math", "20000", "65537"], capture_output=True,
                                   text=True, timeout=300)
                ok = r.returncode == 0
                _libm_state["report"] = r.stdout[-400:]
            except Exception:
                ok = True
        _libm_state["ok"] = ok
    return _libm_state["ok"]


def pytest_collection_modifyitems(config, items):
    gpu_items = [it for it in items if it.get_closest_marker("gpu") is not None
                 and any(k in it.name for k in _EXACT_VS_LIVE_ORACLE)]
    if not gpu_items or os.environ.get("VPT_TESTS_ASSUME_LIBM") == "1":
        return
    if host_libm_is_the_restated_variant():
        return
    skip = pytest.mark.skip(reason="this host's libm is not the glibc FMA+AVX2 variant the device restates "
                                   "(tests/hostsim --selftest-math reports mismatches): the live oracle's bits "
                                   "are not the contract here; the statistical gates apply")
    for it in gpu_items:
        it.add_marker(skip)
