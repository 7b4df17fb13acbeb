"""ctypes binding of include/rtow_b200.h and include/rtow_b200_scenes.h.

Mirrors the reference-facing calls one to one; every function returns through the C status
code and raises RtowError with rtow_last_error() on failure.  There is no Python-side
computation and no fallback path.
"""
import ctypes as C
import os

import numpy as np

from . import _abi as abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class RtowError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"rtow error {code}: {msg}")
        self.code = code


def lib_path():
    # RTOW_B200_LIB: an alternative build of the same library (kernel experiments, tools/build_variant.sh)
    return os.environ.get("RTOW_B200_LIB") or os.path.join(_HERE, "csrc", "librtow_b200.so")


def lib():
    """The CUDA library.  Missing library = hard error (there is no CPU fallback)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback."
        )
    L = C.CDLL(path)
    vp, i32, u64, f32 = C.c_void_p, C.c_int32, C.c_uint64, C.c_float
    P = C.POINTER
    sig = {
        "rtow_device_count": (C.c_int, []),
        "rtow_scene_create": (C.c_int, [P(abi.SceneDesc), C.c_int, P(vp)]),
        "rtow_scene_destroy": (None, [vp]),
        "rtow_trace_rays": (C.c_int, [vp, vp, u64, vp, C.c_int, u64, P(abi.Stats)]),
        "rtow_trace_rays_device": (C.c_int, [vp, vp, u64, vp, C.c_int, u64, vp]),
        "rtow_render": (C.c_int, [vp, P(abi.Camera), P(abi.RenderParams), vp, vp, P(abi.Stats)]),
        "rtow_render_accumulate_device": (C.c_int, [vp, P(abi.Camera), P(abi.RenderParams), vp, vp]),
        "rtow_render_accumulate_device_noalloc": (C.c_int, [vp, P(abi.Camera), P(abi.RenderParams), vp, vp]),
        "rtow_render_accumulate_moments_device": (C.c_int, [vp, P(abi.Camera), P(abi.RenderParams), vp, vp, vp]),
        "rtow_stream_stats": (C.c_int, [vp, vp, P(abi.Stats)]),
        "rtow_resolve_device": (C.c_int, [vp, vp, i32, i32, i32, vp, vp, vp]),
        "rtow_trace_counters": (C.c_int, [vp, vp, u64, C.c_int, P(C.c_double), P(C.c_double)]),
        "rtow_render_counters": (C.c_int, [vp, P(abi.Camera), P(abi.RenderParams), P(abi.WorkCounters)]),
        "rtow_render_profile": (C.c_int, [vp, P(abi.RenderProfile)]),
        "rtow_multi_create": (C.c_int, [P(abi.SceneDesc), P(i32), i32, P(vp)]),
        "rtow_multi_render": (C.c_int, [vp, P(abi.Camera), P(abi.RenderParams), vp, vp, P(abi.Stats)]),
        "rtow_multi_device_count": (C.c_int, [vp]),
        "rtow_multi_peer_mapped": (C.c_int, [vp]),
        "rtow_multi_destroy": (None, [vp]),
        "rtow_last_error": (C.c_char_p, []),
        "rtow_build_info": (C.c_char_p, []),
        "rtow_host_scene_build": (C.c_int, [C.c_char_p, u64, C.c_int, C.c_int, P(vp)]),
        "rtow_host_scene_desc": (P(abi.SceneDesc), [vp]),
        "rtow_host_scene_free": (None, [vp]),
        "rtow_host_set_image": (C.c_int, [vp, i32, i32]),
        "rtow_host_set_image_dir": (C.c_int, [C.c_char_p]),
        "rtow_host_decode_image": (C.c_int, [C.c_char_p, P(P(C.c_uint8)), P(i32), P(i32), P(i32)]),
        "rtow_host_free_image": (None, [P(C.c_uint8)]),
        "rtow_host_camera": (C.c_int, [P(f32 * 3), P(f32 * 3), P(f32 * 3), f32, f32, f32, f32, f32, f32, P(abi.Camera)]),
        "rtow_write_ppm": (C.c_int, [C.c_char_p, vp, i32, i32]),
        "rtow_write_pfm": (C.c_int, [C.c_char_p, vp, i32, i32]),
        "rtow_write_exr": (C.c_int, [C.c_char_p, vp, i32, i32]),
        "rtow_host_camera_oriented": (C.c_int, [P(f32 * 3), P(f32 * 16), f32, f32, f32, f32, f32, f32, P(abi.Camera)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)   # AttributeError here = the library does not export what the header declares
        fn.restype, fn.argtypes = res, args
    _LIB = L
    return L


EXPORTED_SYMBOLS = [
    "rtow_device_count", "rtow_scene_create", "rtow_scene_destroy", "rtow_trace_rays", "rtow_trace_rays_device",
    "rtow_render", "rtow_render_accumulate_device", "rtow_render_accumulate_device_noalloc", "rtow_render_accumulate_moments_device", "rtow_stream_stats", "rtow_resolve_device", "rtow_trace_counters", "rtow_render_counters", "rtow_render_profile",
    "rtow_multi_create", "rtow_multi_render", "rtow_multi_device_count", "rtow_multi_peer_mapped", "rtow_multi_destroy",
    "rtow_last_error", "rtow_build_info", "rtow_host_scene_build", "rtow_host_scene_desc", "rtow_host_scene_free",
    "rtow_host_set_image", "rtow_host_set_image_dir", "rtow_host_decode_image", "rtow_host_free_image", "rtow_host_camera", "rtow_write_ppm", "rtow_write_pfm", "rtow_write_exr", "rtow_host_camera_oriented",
]


def _check(code):
    if code != abi.OK:
        raise RtowError(code, lib().rtow_last_error().decode(errors="replace"))


def _vec3(v):
    return (C.c_float * 3)(*[float(x) for x in v])


def make_camera(look_from, look_at, up, vfov, aspect, aperture, focus, t0=0.0, t1=1.0):
    """Camera(lookFrom, lookAt, up, vFoV, aspect, aperture, focusDistance, t0, t1) -- camera.h:13."""
    cam = abi.Camera()
    _check(lib().rtow_host_camera(C.byref(_vec3(look_from)), C.byref(_vec3(look_at)), C.byref(_vec3(up)),
                                  vfov, aspect, aperture, focus, t0, t1, C.byref(cam)))
    return cam


def oriented_camera(look_from, orientation, vfov, aspect, aperture=2.0, focus=1000.0, t0=0.0, t1=1.0):
    """Camera::setLookFromPoint + setOrientationMatrix (camera.h:166-229): the reference's per-frame fly-camera update.
    orientation: 4x4, rows 0,1,2 = w,u,v."""
    cam = abi.Camera()
    m = (C.c_float * 16)(*[float(x) for x in np.asarray(orientation, dtype=np.float32).reshape(16)])
    _check(lib().rtow_host_camera_oriented(C.byref(_vec3(look_from)), C.byref(m), vfov, aspect, aperture, focus, t0, t1, C.byref(cam)))
    return cam


def main_camera(width, height):
    """The camera main() builds: main.cpp:111-125."""
    return make_camera((0, 0, 0), (1, 0, 0), (0, 0, -1), 60.0, float(width) / float(height), 2.0, 1000.0, 0.0, 1.0)


def image_dir():
    """Where the builders' ./input_images/*.jpg (scenes.h:47,140) are looked for when the CWD does not have them:
    $RTOW_INPUT_IMAGES, else the copies build() stages next to the library (git-ignored, they travel with the snapshot),
    else the reference checkout."""
    for d in (os.environ.get("RTOW_INPUT_IMAGES"), os.path.join(_HERE, "input_images"),
              "/root/reference/rayTracingOneWeekend/input_images"):
        if d and os.path.isfile(os.path.join(d, "1_earth_8k.jpg")):
            return d
    return None


def decode_image(path):
    """The library's own JPEG / PPM decoder (include/rtow/jpeg_decode.h): texels as stbi_load returns them."""
    L = lib()
    px, nx, ny, nc = C.POINTER(C.c_uint8)(), C.c_int32(), C.c_int32(), C.c_int32()
    _check(L.rtow_host_decode_image(str(path).encode(), C.byref(px), C.byref(nx), C.byref(ny), C.byref(nc)))
    try:
        return np.ctypeslib.as_array(px, shape=(ny.value, nx.value, nc.value)).copy()
    finally:
        L.rtow_host_free_image(px)


class HostScene:
    """A scene built by the host-side scene API (mirrors scenes.h) and flattened.

    image=None: the builders load ./input_images/*.jpg themselves like the reference (decoded by the library);
    image=array: those loads return the given texels; image=False: loads fail, i.e. the reference's glass-sphere
    fallback (scenes.h:104-109,202-207)."""

    def __init__(self, name, seed=12345, n=0, wrap=True, image=None):
        L = lib()
        d = image_dir() if image is None else None
        _check(L.rtow_host_set_image_dir(d.encode() if d else (b"/nonexistent" if image is False else None)))
        if image is False:
            image = None
        if image is not None:
            img = np.ascontiguousarray(image, dtype=np.uint8)
            assert img.ndim == 3 and img.shape[2] == 3
            _check(L.rtow_host_set_image(img.ctypes.data, img.shape[1], img.shape[0]))
        else:
            _check(L.rtow_host_set_image(None, 0, 0))
        h = C.c_void_p()
        try:
            _check(L.rtow_host_scene_build(name.encode(), seed, n, 1 if wrap else 0, C.byref(h)))
        finally:
            L.rtow_host_set_image(None, 0, 0)
        self._h = h
        self.name = name

    @property
    def desc(self):
        return lib().rtow_host_scene_desc(self._h).contents

    def arrays(self):
        return abi.desc_arrays(self.desc)

    def close(self):
        if self._h:
            lib().rtow_host_scene_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceScene:
    """rtow_scene*: device copies of a flattened scene + the traversal structures."""

    def __init__(self, desc, device=0):
        L = lib()
        h = C.c_void_p()
        _check(L.rtow_scene_create(C.byref(desc), device, C.byref(h)))
        self._h = h
        self.device = device

    def close(self):
        if self._h:
            lib().rtow_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- world->hit() ---------------------------------------------------------------------
    def trace(self, rays, traverse=abi.TRAVERSE_AUTO, seed=0):
        rays = np.ascontiguousarray(rays, dtype=abi.ray_dtype)
        hits = np.zeros(len(rays), dtype=abi.hit_dtype)
        st = abi.Stats()
        _check(lib().rtow_trace_rays(self._h, rays.ctypes.data, len(rays), hits.ctypes.data, traverse, seed, C.byref(st)))
        return hits, st

    def trace_device(self, d_rays_ptr, n, d_hits_ptr, traverse=abi.TRAVERSE_AUTO, seed=0, stream=None):
        _check(lib().rtow_trace_rays_device(self._h, d_rays_ptr, n, d_hits_ptr, traverse, seed, stream))

    def trace_counters(self, rays, traverse=abi.TRAVERSE_AUTO):
        rays = np.ascontiguousarray(rays, dtype=abi.ray_dtype)
        nv, pt = C.c_double(), C.c_double()
        _check(lib().rtow_trace_counters(self._h, rays.ctypes.data, len(rays), traverse, C.byref(nv), C.byref(pt)))
        return nv.value, pt.value

    # -- the pixel loop -------------------------------------------------------------------
    def render(self, cam, params, want_bgra=False):
        W, H = params.width, (params.row_count if params.row_count > 0 else params.height)    # a band of rows gives band-sized outputs
        rgb = np.zeros((H, W, 3), dtype=np.float32)
        bgra = np.zeros((H, W, 4), dtype=np.uint8) if want_bgra else None
        st = abi.Stats()
        _check(lib().rtow_render(self._h, C.byref(cam), C.byref(params), rgb.ctypes.data,
                                 bgra.ctypes.data if want_bgra else None, C.byref(st)))
        return (rgb, bgra, st) if want_bgra else (rgb, st)

    def render_into(self, cam, params, rgb_out, bgra_out=None):
        """Host buffers supplied by the caller (pinned buffers in bench.py's e2e leg)."""
        st = abi.Stats()
        _check(lib().rtow_render(self._h, C.byref(cam), C.byref(params), rgb_out, bgra_out, C.byref(st)))
        return st

    def render_counters(self, cam, params):
        out = abi.WorkCounters()
        _check(lib().rtow_render_counters(self._h, C.byref(cam), C.byref(params), C.byref(out)))
        return out

    def render_profile(self):
        out = abi.RenderProfile()
        _check(lib().rtow_render_profile(self._h, C.byref(out)))
        return out

    def accumulate_device(self, cam, params, d_accum_ptr, stream=None):
        _check(lib().rtow_render_accumulate_device(self._h, C.byref(cam), C.byref(params), d_accum_ptr, stream))

    def accumulate_device_noalloc(self, cam, params, d_accum_ptr, stream=None):
        """Static-scratch variant: no device allocation at call time (include/rtow_b200.h)."""
        _check(lib().rtow_render_accumulate_device_noalloc(self._h, C.byref(cam), C.byref(params), d_accum_ptr, stream))

    def accumulate_moments_device(self, cam, params, d_accum_ptr, d_accum_sq_ptr, stream=None):
        _check(lib().rtow_render_accumulate_moments_device(self._h, C.byref(cam), C.byref(params), d_accum_ptr, d_accum_sq_ptr, stream))

    def stream_stats(self, stream=None):
        st = abi.Stats()
        _check(lib().rtow_stream_stats(self._h, stream, C.byref(st)))
        return st

    def resolve_device(self, d_accum_ptr, width, height, spp, d_rgb_ptr=None, d_bgra_ptr=None, stream=None):
        _check(lib().rtow_resolve_device(self._h, d_accum_ptr, width, height, spp, d_rgb_ptr, d_bgra_ptr, stream))


class MultiDeviceScene:
    """rtow_multi*: the scene on several GPUs of this box, one frame rendered by all of them (spp-sharded) from this process."""

    def __init__(self, desc, devices=None):
        L = lib()
        h = C.c_void_p()
        if devices is None:
            _check(L.rtow_multi_create(C.byref(desc), None, 0, C.byref(h)))
        else:
            arr = (C.c_int32 * len(devices))(*devices)
            _check(L.rtow_multi_create(C.byref(desc), arr, len(devices), C.byref(h)))
        self._h = h
        self.shards = L.rtow_multi_device_count(h)
        self.peer_mapped = L.rtow_multi_peer_mapped(h)

    def render(self, cam, params, want_bgra=False):
        W, H = params.width, (params.row_count if params.row_count > 0 else params.height)
        rgb = np.zeros((H, W, 3), dtype=np.float32)
        bgra = np.zeros((H, W, 4), dtype=np.uint8) if want_bgra else None
        st = abi.Stats()
        _check(lib().rtow_multi_render(self._h, C.byref(cam), C.byref(params), rgb.ctypes.data, bgra.ctypes.data if want_bgra else None, C.byref(st)))
        return (rgb, bgra, st) if want_bgra else (rgb, st)

    def render_into(self, cam, params, rgb_out, bgra_out=None):
        st = abi.Stats()
        _check(lib().rtow_multi_render(self._h, C.byref(cam), C.byref(params), rgb_out, bgra_out, C.byref(st)))
        return st

    def close(self):
        if self._h:
            lib().rtow_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def write_ppm(path, rgb):
    rgb = np.ascontiguousarray(rgb, dtype=np.float32)
    _check(lib().rtow_write_ppm(str(path).encode(), rgb.ctypes.data, rgb.shape[1], rgb.shape[0]))


def write_pfm(path, rgb):
    rgb = np.ascontiguousarray(rgb, dtype=np.float32)
    _check(lib().rtow_write_pfm(str(path).encode(), rgb.ctypes.data, rgb.shape[1], rgb.shape[0]))


def write_exr(path, rgb):
    rgb = np.ascontiguousarray(rgb, dtype=np.float32)
    _check(lib().rtow_write_exr(str(path).encode(), rgb.ctypes.data, rgb.shape[1], rgb.shape[0]))
