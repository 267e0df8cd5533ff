"""GPU: SURVEY 8(f) #3 - tesselate_surfaces with Catmull-Clark subdivision, smooth normals and
texture displacement on the device (csrc/tesselate.cuh, host/vpt_tesselate.cpp) against the
reference's own tessellation (tools/vpt_export runs yocto's tesselate_surfaces when it writes
scenes/_built/<scene>.vptscene; <scene>.subdivs.vptarch holds the control cages): positions,
normals, texture coordinates and triangles of every subdivided shape, bit for bit."""
import os
import subprocess

import numpy as np
import pytest

import util

pytestmark = pytest.mark.gpu
CLI = os.path.join(util.ROOT, "volumetric-path-tracer_b200", "ypathtrace")


@pytest.mark.parametrize("scene", ["01_surface", "02_rollingteapot"])
def test_device_tessellation_equals_the_reference(vpt, scene):
    path = util.scene_path(scene)
    cages = os.path.join(util.BUILT, scene + ".subdivs.vptarch")
    if path is None or not os.path.exists(cages):
        pytest.skip(f"{scene} or its control cages not built")
    ar = vpt.archive.load(cages)
    n = int(ar["subdivs.count"][0])
    assert n >= 1
    shapes = [int(ar[f"subdiv{i}.params"][4]) for i in range(n)]
    ref = vpt.binding.Scene(path)   # tessellated by the reference when the archive was written
    dev = vpt.binding.Scene(path)
    dev.load_subdivs(cages)
    dev.tesselate()                 # those shapes are emptied and rebuilt on the device
    displaced = 0
    for i, shape in enumerate(shapes):
        a, b = ref.shape(shape), dev.shape(shape)
        levels, _, smooth, disp_tex, _ = (int(v) for v in ar[f"subdiv{i}.params"])
        displaced += int(disp_tex >= 0 and float(ar[f"subdiv{i}.displacement"][0]) != 0)
        assert len(a["triangles"]) > len(ar[f"subdiv{i}.quadspos"]), "the cage should have been subdivided"
        assert np.array_equal(a["triangles"], b["triangles"]), f"subdiv {i}: triangles"
        for key in ("positions", "normals", "texcoords"):
            assert a[key].shape == b[key].shape, f"subdiv {i}: {key} count"
            assert np.array_equal(a[key].view(np.uint32), b[key].view(np.uint32)), \
                f"subdiv {i} ({levels} levels, smooth {smooth}): {key} differ"
        assert len(b["quads"]) == 0
    if scene == "01_surface":
        assert displaced >= 1, "01_surface has a displaced subdiv (sphere-displaced)"
    ref.close(), dev.close()


def test_cli_tessellates_kept_subdivs(vpt, tmp_path):
    """An archive exported with --keep-subdivs renders (ypathtrace tessellates on the device) to the
    same float image as the archive the reference tessellated."""
    a = util.scene_path("02_rollingteapot")
    b = util.scene_path("02_rollingteapot_subdiv")
    if a is None or b is None:
        pytest.skip("scenes not built")
    outs = []
    for path in (a, b):
        out = str(tmp_path / (os.path.basename(path) + ".pfm"))
        c = subprocess.run([CLI, "--scene", path, "--output", out, "--shader", "pathtrace", "--samples", "2",
                            "--resolution", "160"], capture_output=True, text=True)
        assert c.returncode == 0, c.stdout + c.stderr
        outs.append((open(out, "rb").read(), c.stdout))
    assert "tesselate surfaces (device)" in outs[1][1] and "tesselate surfaces" not in outs[0][1]
    assert outs[0][0] == outs[1][0]


def test_catmullclark_entry_point_small_cases(vpt):
    """vpt_cuda_catmullclark through the C ABI on a closed cube and on an open strip with a
    triangle (boundary + z == w paths): sizes follow V' = V + E + F, quads' = sum of face arity."""
    import ctypes as C

    cuda, _ = vpt.binding.libs()
    vpt.binding.set_device(0)
    cube_q = np.array([[0, 1, 2, 3], [4, 7, 6, 5], [0, 4, 5, 1], [1, 5, 6, 2], [2, 6, 7, 3], [3, 7, 4, 0]], np.int32)
    cube_v = np.array([[-1, -1, -1], [1, -1, -1], [1, 1, -1], [-1, 1, -1], [-1, -1, 1], [1, -1, 1], [1, 1, 1],
                       [-1, 1, 1]], np.float32)
    strip_q = np.array([[0, 1, 4, 3], [1, 2, 5, 4], [3, 4, 6, 6]], np.int32)
    strip_v = np.array([[0, 0], [1, 0], [2, 0], [0, 1], [1, 1], [2, 1], [0.5, 2]], np.float32)
    for quads, verts, lock, edges in ((cube_q, cube_v, 0, 12), (strip_q, strip_v, 1, 9)):
        oq, ov = C.POINTER(C.c_int32)(), C.POINTER(C.c_float)()
        nq, nv = C.c_int32(), C.c_int32()
        dim = verts.shape[1]
        assert cuda.vpt_cuda_catmullclark(quads.ctypes.data_as(C.POINTER(C.c_int32)), len(quads),
                                          verts.ctypes.data_as(C.POINTER(C.c_float)), len(verts), dim, lock, 1,
                                          C.byref(oq), C.byref(nq), C.byref(ov), C.byref(nv)) == 0
        arity = sum(3 if q[2] == q[3] else 4 for q in quads)
        assert nq.value == arity and nv.value == len(verts) + edges + len(quads)
        out_v = np.ctypeslib.as_array(ov, shape=(nv.value, dim)).copy()
        out_q = np.ctypeslib.as_array(oq, shape=(nq.value, 4)).copy()
        cuda.vpt_cuda_free_host.argtypes = [C.c_void_p]
        cuda.vpt_cuda_free_host(oq), cuda.vpt_cuda_free_host(ov)
        assert np.isfinite(out_v).all() and out_q.min() >= 0 and out_q.max() < nv.value
        if lock:  # locked boundary: the original boundary vertices stay where they were
            assert np.array_equal(out_v[[0, 2, 6]], strip_v[[0, 2, 6]])
        else:     # a cube shrinks towards the sphere, symmetric in all corners
            assert np.allclose(np.abs(out_v[:8]), np.abs(out_v[0]))
