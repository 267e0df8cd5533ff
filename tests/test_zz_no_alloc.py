"""GPU: vpt_cuda_intersect_noalloc — the allocation-free variant of vpt_cuda_intersect (static
__device__ staging array inside libvpt_cuda.so, include/vpt_cuda.h). It has no counterpart in
the reference; it was added for callers that cannot call
cudaMalloc. Called here straight through the C ABI of libvpt_cuda.so and compared with the
existing entry point bit for bit (integer ids and float bits: the tolerance every first-hit
test of this suite uses), on two calls in a row that share the array.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu

MAX_RAYS = 65536  # VPT_NOALLOC_MAX_RAYS


@pytest.fixture()
def abi(vpt):
    cuda, host = vpt.binding.libs()
    host.vpth_scene_device_handle.restype = C.c_void_p
    host.vpth_scene_device_handle.argtypes = [C.c_void_p]
    for fn in (cuda.vpt_cuda_intersect, cuda.vpt_cuda_intersect_noalloc):
        fn.restype = C.c_int
        fn.argtypes = [C.c_void_p, C.POINTER(C.c_float), C.c_int, C.c_int, C.POINTER(C.c_int32),
                       C.POINTER(C.c_float)]
    vpt.binding.set_device(0)
    scene = vpt.binding.Scene(util.scene_path("mixed"))
    handle = host.vpth_scene_device_handle(scene._h)
    assert handle, "scene upload failed"
    yield cuda, handle
    scene.close()


def query(fn, handle, rays, instance=-1):
    rays = np.ascontiguousarray(rays, dtype=np.float32)
    n = rays.shape[0]
    # poison the outputs: every record must be written by the call
    ids = np.full((n, 3), -77, np.int32)
    uvd = np.full((n, 3), np.nan, np.float32)
    status = fn(handle, rays.ctypes.data_as(C.POINTER(C.c_float)), n, instance,
                ids.ctypes.data_as(C.POINTER(C.c_int32)), uvd.ctypes.data_as(C.POINTER(C.c_float)))
    return status, ids, uvd


def test_noalloc_intersect_equals_intersect_on_two_calls_in_a_row(abi):
    cuda, handle = abi
    first = util.camera_rays(0, 0, np.random.default_rng(31), 50000)
    second = util.camera_rays(0, 0, np.random.default_rng(32), 20000)  # shorter: stale tail behind it
    for rays, instance in ((first, -1), (second, -1), (second, 9), (first[:1], -1)):
        s0, ids0, uvd0 = query(cuda.vpt_cuda_intersect, handle, rays, instance)
        s1, ids1, uvd1 = query(cuda.vpt_cuda_intersect_noalloc, handle, rays, instance)
        assert s0 == 0 and s1 == 0, cuda.vpt_cuda_last_error()
        assert np.array_equal(ids1, ids0)
        assert not util.bit_mismatch(uvd1, uvd0).any()
        assert (ids1 != -77).all() and not np.isnan(uvd1).any()
    assert ids0.shape == (1, 3)
    s, ids, _ = query(cuda.vpt_cuda_intersect_noalloc, handle, first, -1)
    assert s == 0 and ids[:, 0].sum() > 1000, "ray batch should hit the scene"


def test_noalloc_intersect_limit_and_errors(abi):
    cuda, handle = abi
    full = util.camera_rays(0, 0, np.random.default_rng(33), MAX_RAYS)
    s0, ids0, uvd0 = query(cuda.vpt_cuda_intersect, handle, full)
    s1, ids1, uvd1 = query(cuda.vpt_cuda_intersect_noalloc, handle, full)  # exactly the maximum
    assert s0 == 0 and s1 == 0
    assert np.array_equal(ids1, ids0) and not util.bit_mismatch(uvd1, uvd0).any()
    over = np.concatenate([full, full[:1]])
    s, ids, _ = query(cuda.vpt_cuda_intersect_noalloc, handle, over)
    assert s == -1  # VPT_ERR_INVALID
    msg = cuda.vpt_cuda_last_error().decode()
    assert "65536" in msg and "exceed" in msg
    assert (ids == -77).all(), "a rejected call must not touch the outputs"
    s, _, _ = query(cuda.vpt_cuda_intersect_noalloc, handle, full[:4], instance=10 ** 6)
    assert s == -1
    s, ids, _ = query(cuda.vpt_cuda_intersect_noalloc, handle, full[:0])  # empty batch
    assert s == 0 and ids.shape == (0, 3)
