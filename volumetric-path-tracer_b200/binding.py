"""ctypes binding of the C ABI (include/vpt_cuda.h, include/vpt_host.h).

Python is test / bench plumbing only: every call below lands in libvpt_host.so
(the C++ facade mirroring yocto_pathtrace) and from there in libvpt_cuda.so
(the sm_100a kernels). There is no Python or CPU rendering path; a missing
library raises at import time and a missing GPU raises on the first compute call.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# VPT_LIB_DIR: load another build of the two libraries (kernel experiments); default in-tree
_LIB_DIR = os.environ.get("VPT_LIB_DIR") or _HERE
LIB_CUDA_PATH = os.path.join(_LIB_DIR, "libvpt_cuda.so")
LIB_HOST_PATH = os.path.join(_LIB_DIR, "libvpt_host.so")

SHADERS = ["volpathtrace", "pathtrace", "naive", "eyelight", "normal", "texcoord", "color",
           "implicit", "implicit_normal"]
PIPELINES = {"auto": 0, "mega": 1, "wavefront": 2}

VPT_OK, VPT_ERR_INVALID, VPT_ERR_UNSUPPORTED, VPT_ERR_NO_DEVICE, VPT_ERR_CUDA = 0, -1, -2, -3, -4


class VptError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"vpt status {status}: {message}")
        self.status = status


class Params(C.Structure):
    """vpt_params (pathtrace_params, yocto_pathtrace.h:87-99, as plain C)."""
    _fields_ = [("camera", C.c_int32), ("resolution", C.c_int32), ("shader", C.c_int32),
                ("samples", C.c_int32), ("bounces", C.c_int32), ("noparallel", C.c_int32),
                ("pratio", C.c_int32), ("exposure", C.c_float), ("filmic", C.c_int32),
                ("noimplicit_mis", C.c_int32), ("spheretrace_maxiter", C.c_int32),
                ("pipeline", C.c_int32)]


class TraceStats(C.Structure):
    _fields_ = [("device_ms", C.c_float), ("top_kernel_ms", C.c_float),
                ("kernel_launches", C.c_int32), ("top_kernel_launches", C.c_int32),
                ("paths", C.c_uint64), ("rays_extend", C.c_uint64),
                ("rays_lightpdf", C.c_uint64), ("sdf_evals", C.c_uint64),
                ("top_kernel_units", C.c_uint64), ("graph_launches", C.c_uint64)]


class BvhBuildStats(C.Structure):
    _fields_ = [("primitives", C.c_int32), ("nodes", C.c_int32), ("keys_ms", C.c_float),
                ("sort_ms", C.c_float), ("hierarchy_ms", C.c_float)]


class LightsStats(C.Structure):
    _fields_ = [("num_lights", C.c_int32), ("cdf_elements", C.c_uint64),
                ("weights_ms", C.c_float), ("scan_ms", C.c_float)]


BVH_NODE_DTYPE = np.dtype([("bbox_min", np.float32, 3), ("bbox_max", np.float32, 3),
                           ("start", np.int32), ("num", np.int16), ("axis", np.int8),
                           ("internal", np.uint8)])
assert BVH_NODE_DTYPE.itemsize == 32


def make_params(shader="pathtrace", samples=512, resolution=720, bounces=4, camera=0,
                noimplicit_mis=False, spheretrace_maxiter=450, pipeline="auto") -> Params:
    p = Params()
    p.camera, p.resolution, p.shader = camera, resolution, SHADERS.index(shader)
    p.samples, p.bounces, p.noparallel, p.pratio = samples, bounces, 0, 8
    p.exposure, p.filmic = 0.0, 0
    p.noimplicit_mis, p.spheretrace_maxiter = int(noimplicit_mis), spheretrace_maxiter
    p.pipeline = PIPELINES[pipeline]
    return p


_libs = None


def libs():
    """Loads libvpt_cuda.so then libvpt_host.so; raises if either is missing."""
    global _libs
    if _libs is None:
        for path in (LIB_CUDA_PATH, LIB_HOST_PATH):
            if not os.path.exists(path):
                raise ImportError(f"{path} is missing: build it with __graft_entry__.build(); "
                                  "there is no fallback implementation")
        cuda = C.CDLL(LIB_CUDA_PATH, mode=C.RTLD_GLOBAL)
        host = C.CDLL(LIB_HOST_PATH, mode=C.RTLD_GLOBAL)
        cuda.vpt_cuda_last_error.restype = C.c_char_p
        host.vpth_last_error.restype = C.c_char_p
        host.vpth_render_device_image.restype = C.c_void_p
        cuda.vpt_cuda_image_device_ptr.restype = C.c_void_p
        cuda.vpt_cuda_scene_bytes.restype = C.c_size_t
        _libs = (cuda, host)
    return _libs


def device_count() -> int:
    return libs()[0].vpt_cuda_device_count()


def set_device(device: int) -> None:
    cuda = libs()[0]
    status = cuda.vpt_cuda_set_device(device)
    if status != VPT_OK:
        raise VptError(status, cuda.vpt_cuda_last_error().decode())


def _check(status: int) -> None:
    if status != VPT_OK:
        cuda, host = libs()
        msg = host.vpth_last_error().decode() or cuda.vpt_cuda_last_error().decode()
        raise VptError(status, msg)


def _fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _iptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


class Scene:
    """load_scene + make_bvh + make_lights on the host; device upload is lazy."""

    def __init__(self, path: str):
        self._host = libs()[1]
        self._h = C.c_void_p()
        _check(self._host.vpth_load_scene(path.encode(), C.byref(self._h)))
        self.path = path

    def close(self):
        if self._h:
            self._host.vpth_free_scene(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def counts(self) -> dict:
        v = (C.c_int32 * 12)()
        _check(self._host.vpth_scene_counts(self._h, v))
        names = ["cameras", "instances", "shapes", "materials", "textures", "environments",
                 "volumes", "vol_instances", "sdfs", "lights", "scene_nodes", "shape_nodes"]
        return dict(zip(names, list(v)))

    def bvh(self, shape: int = -1):
        """(nodes, primitives) of one shape's BVH, or of the scene BVH for shape < 0."""
        nn, npr = C.c_int32(), C.c_int32()
        _check(self._host.vpth_bvh_size(self._h, shape, C.byref(nn), C.byref(npr)))
        nodes = np.zeros(nn.value, dtype=BVH_NODE_DTYPE)
        prims = np.zeros(npr.value, dtype=np.int32)
        _check(self._host.vpth_bvh_copy(self._h, shape, nodes.ctypes.data_as(C.c_void_p),
                                        _iptr(prims)))
        return nodes, prims

    def lights(self):
        out = []
        for i in range(self.counts()["lights"]):
            ids, n = (C.c_int32 * 3)(), C.c_int32()
            _check(self._host.vpth_light_info(self._h, i, ids, C.byref(n)))
            cdf = np.zeros(n.value, dtype=np.float32)
            if n.value:
                _check(self._host.vpth_light_cdf(self._h, i, _fptr(cdf)))
            out.append({"instance": ids[0], "environment": ids[1], "sdf": ids[2], "cdf": cdf})
        return out

    def use_device_lights(self) -> dict:
        """make_lights on the device; returns the construction timings."""
        st = LightsStats()
        _check(self._host.vpth_scene_use_device_lights(self._h, C.byref(st)))
        return {"lights": st.num_lights, "cdf_elements": st.cdf_elements,
                "weights_ms": st.weights_ms, "scan_ms": st.scan_ms}

    def device_lights(self):
        """The light list and CDFs as they lie in HBM."""
        n = C.c_int32()
        _check(self._host.vpth_device_light(self._h, -1, None, None, C.c_size_t(0), C.byref(n)))
        out = []
        for i in range(n.value):
            ids, m = (C.c_int32 * 3)(), C.c_int32()
            _check(self._host.vpth_device_light(self._h, i, ids, None, C.c_size_t(0), C.byref(m)))
            cdf = np.zeros(m.value, dtype=np.float32)
            if m.value:
                _check(self._host.vpth_device_light(self._h, i, ids, _fptr(cdf), C.c_size_t(cdf.size), C.byref(m)))
            out.append({"instance": ids[0], "environment": ids[1], "sdf": ids[2], "cdf": cdf})
        return out

    def load_subdivs(self, path: str) -> None:
        _check(self._host.vpth_scene_load_subdivs(self._h, path.encode()))

    def tesselate(self) -> None:
        """tesselate_surfaces with Catmull-Clark / normals / displacement on the device."""
        _check(self._host.vpth_scene_tesselate(self._h))

    def shape(self, index: int) -> dict:
        sizes = (C.c_int32 * 6)()
        _check(self._host.vpth_shape_sizes(self._h, index, sizes))
        out = {"positions": np.zeros((sizes[0], 3), np.float32), "normals": np.zeros((sizes[1], 3), np.float32),
               "texcoords": np.zeros((sizes[2], 2), np.float32), "triangles": np.zeros((sizes[3], 3), np.int32),
               "quads": np.zeros((sizes[4], 4), np.int32)}
        _check(self._host.vpth_shape_copy(self._h, index, _fptr(out["positions"]), _fptr(out["normals"]),
                                          _fptr(out["texcoords"]), _iptr(out["triangles"]), _iptr(out["quads"])))
        return out

    def use_device_bvh(self) -> BvhBuildStats:
        """Rebuild the shape BVHs on the device (LBVH, vpt_cuda_build_shape_bvh)."""
        st = BvhBuildStats()
        _check(self._host.vpth_scene_use_device_bvh(self._h, C.byref(st)))
        return st

    def intersect(self, rays: np.ndarray, instance: int = -1):
        """rays: [n, 8] float32 -> (ids [n,3] int32: hit, instance, element; uvd [n,3])."""
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = rays.shape[0]
        ids = np.zeros((n, 3), dtype=np.int32)
        uvd = np.zeros((n, 3), dtype=np.float32)
        _check(self._host.vpth_intersect(self._h, _fptr(rays), n, instance, _iptr(ids), _fptr(uvd)))
        return ids, uvd

    def spheretrace(self, rays: np.ndarray, maxiter: int = 450):
        rays = np.ascontiguousarray(rays, dtype=np.float32)
        n = rays.shape[0]
        ids = np.zeros((n, 3), dtype=np.int32)
        dn = np.zeros((n, 4), dtype=np.float32)
        _check(self._host.vpth_spheretrace(self._h, _fptr(rays), n, maxiter, _iptr(ids), _fptr(dn)))
        return ids, dn


class Render:
    """make_state + pathtrace_samples + get_render for one scene / params pair."""

    def __init__(self, scene: Scene, params: Params, gpus: int = 1):
        self._host = libs()[1]
        self.scene, self.params = scene, params
        self._h = C.c_void_p()
        _check(self._host.vpth_render_create(scene._h, C.byref(params), C.byref(self._h)))
        if gpus > 1:  # sample-range sharding over GPUs of this box, summed inside image()
            _check(self._host.vpth_render_set_gpus(self._h, gpus))
        w, h = C.c_int32(), C.c_int32()
        _check(self._host.vpth_render_size(self._h, C.byref(w), C.byref(h)))
        self.width, self.height = w.value, h.value

    def close(self):
        if self._h:
            self._host.vpth_render_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def rngs(self) -> np.ndarray:
        out = np.zeros((self.width * self.height, 2), dtype=np.uint64)
        _check(self._host.vpth_render_get_rngs(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def pin(self):
        """Page-lock the host mirrors now (default: after the third host transfer)."""
        _check(self._host.vpth_render_pin(self._h))

    def reset(self, rng_offset: int = 0) -> None:
        _check(self._host.vpth_render_reset(self._h, C.c_uint64(rng_offset)))

    def trace(self, nsamples: int) -> None:
        _check(self._host.vpth_render_samples(self._h, nsamples))

    @property
    def samples(self) -> int:
        return self._host.vpth_render_done_samples(self._h)

    def image(self, normalized: bool = True) -> np.ndarray:
        out = np.empty((self.height, self.width, 4), dtype=np.float32)
        _check(self._host.vpth_render_get_image(self._h, _fptr(out), int(normalized)))
        return out

    def image_view(self) -> np.ndarray:
        """Un-normalised accumulation image [h, w, 4] as a read-only VIEW of the render's host mirror
        (no copy, no allocation): valid until the next call on this render."""
        ptr = C.POINTER(C.c_float)()
        _check(self._host.vpth_render_image_view(self._h, C.byref(ptr)))
        arr = np.ctypeslib.as_array(ptr, shape=(self.height, self.width, 4))
        arr.flags.writeable = False
        return arr

    def image_srgb8(self) -> np.ndarray:
        """Output stage on the device: get_render + rgb_to_srgb + float_to_byte -> [h, w, 4] uint8."""
        out = np.zeros((self.height, self.width, 4), dtype=np.uint8)
        _check(self._host.vpth_render_get_srgb8(self._h, out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def image_tonemapped8(self, exposure: float = 0.0, filmic: bool = False) -> np.ndarray:
        """The viewer's display conversion on the device: tonemap(exposure, filmic) + 8 bit."""
        out = np.zeros((self.height, self.width, 4), dtype=np.uint8)
        _check(self._host.vpth_render_get_tonemapped8(self._h, C.c_float(exposure), int(filmic),
                                                      out.ctypes.data_as(C.POINTER(C.c_uint8))))
        return out

    def stats(self) -> TraceStats:
        s = TraceStats()
        _check(self._host.vpth_render_stats(self._h, C.byref(s)))
        return s

    def export_image(self, device_ptr: int) -> None:
        _check(self._host.vpth_render_export_image(self._h, C.c_void_p(device_ptr)))

    def import_image(self, device_ptr: int, total_samples: int) -> None:
        _check(self._host.vpth_render_import_image(self._h, C.c_void_p(device_ptr), total_samples))

    def device_image_ptr(self) -> int:
        return self._host.vpth_render_device_image(self._h) or 0


def render(scene_path: str, shader: str, samples: int, resolution: int, bounces: int = 4,
           pipeline: str = "auto", **kw) -> np.ndarray:
    """One-call convenience: the un-normalised accumulation image [h, w, 4]."""
    sc = Scene(scene_path)
    r = Render(sc, make_params(shader=shader, samples=samples, resolution=resolution,
                               bounces=bounces, pipeline=pipeline, **kw))
    r.trace(samples)
    img = r.image(normalized=False)
    r.close()
    sc.close()
    return img
