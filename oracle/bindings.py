"""ctypes bindings of the two CPU checkers.  TEST INFRASTRUCTURE ONLY.

  RefOracle   oracle/_ref/libref_oracle.so -- the reference's own code (built from /root/reference
              by oracle/Makefile; prebuilt file travels to the GPU box).
  PortOracle  oracle/_build/librtow_oracle.so -- the plain-C restatement (rtow_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The product package never does.
"""
import ctypes as C
import importlib.util
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)


def load_package():
    """The product's Python package (hyphenated directory) under the module name rtow_b200."""
    if "rtow_b200" in sys.modules:
        return sys.modules["rtow_b200"]
    pkg_dir = os.path.join(_ROOT, "ray-tracing-one-weekend_b200")
    spec = importlib.util.spec_from_file_location(
        "rtow_b200", os.path.join(pkg_dir, "__init__.py"), submodule_search_locations=[pkg_dir]
    )
    mod = importlib.util.module_from_spec(spec)
    sys.modules["rtow_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


abi = load_package().abi

REF_SO = os.path.join(_HERE, "_ref", "libref_oracle.so")
REF_DOF_SO = os.path.join(_HERE, "_ref", "libref_oracle_dof.so")      # the reference built with CAMERA_DOF_EN 1
PORT_SO = os.path.join(_HERE, "_build", "librtow_oracle.so")
REF_IMAGE_DIR = os.path.join(_HERE, "_ref")


def build(verbose=False):
    """make -C oracle: the C port always; the reference bridge when /root/reference exists."""
    out = subprocess.run(["make", "-C", _HERE, "all"], capture_output=True, text=True)
    if verbose or out.returncode != 0:
        sys.stderr.write(out.stdout + out.stderr)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed")


def have_ref():
    return os.path.exists(REF_SO)


def have_ref_dof():
    return os.path.exists(REF_DOF_SO)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class RefScene:
    def __init__(self, oracle, handle, name):
        self.o, self.h, self.name = oracle, handle, name

    @property
    def desc(self):
        d = self.o.L.ref_scene_desc(self.h)
        if not d:
            raise RuntimeError("reference flatten failed: " + self.o.L.ref_last_error().decode())
        return d.contents

    def arrays(self):
        return abi.desc_arrays(self.desc)

    def trace(self, rays, seed=0, threads=1):
        rays = np.ascontiguousarray(rays, dtype=abi.ray_dtype)
        hits = np.zeros(len(rays), dtype=abi.hit_dtype)
        secs = self.o.L.ref_trace(self.h, _ptr(rays), len(rays), _ptr(hits), seed, threads)
        return hits, secs

    def render(self, cam_handle, W, H, spp, seed=1, threads=1, want_moments=False):
        mean = np.zeros((H, W, 3), dtype=np.float32)
        sm = np.zeros((H, W, 3), dtype=np.float64) if want_moments else None
        sq = np.zeros((H, W, 3), dtype=np.float64) if want_moments else None
        rays, paths = C.c_uint64(), C.c_uint64()
        secs = self.o.L.ref_render(self.h, cam_handle, W, H, spp, seed, threads, _ptr(mean),
                                   _ptr(sm) if want_moments else None, _ptr(sq) if want_moments else None,
                                   C.byref(rays), C.byref(paths))
        return dict(mean=mean, sum=sm, sumsq=sq, rays=rays.value, paths=paths.value, seconds=secs)

    def color(self, rays, seed=0):
        rays = np.ascontiguousarray(rays, dtype=abi.ray_dtype)
        rgb = np.zeros((len(rays), 3), dtype=np.float32)
        self.o.L.ref_color(self.h, _ptr(rays), len(rays), seed, _ptr(rgb))
        return rgb

    def texture_value(self, tex_id, uv, p):
        uv = np.ascontiguousarray(uv, dtype=np.float32)
        p = np.ascontiguousarray(p, dtype=np.float32)
        rgb = np.zeros((len(p), 3), dtype=np.float32)
        rc = self.o.L.ref_texture_value(self.h, tex_id, _ptr(uv), _ptr(p), len(p), _ptr(rgb))
        if rc:
            raise RuntimeError("ref_texture_value failed")
        return rgb

    def scatter(self, mat_id, ray, hit, seed):
        atten = (C.c_float * 3)()
        em = (C.c_float * 3)()
        out = np.zeros(1, dtype=abi.ray_dtype)
        ray = np.ascontiguousarray(ray, dtype=abi.ray_dtype)
        hit = np.ascontiguousarray(hit, dtype=abi.hit_dtype)
        ok = self.o.L.ref_material_scatter(self.h, mat_id, _ptr(ray), _ptr(hit), seed, atten, _ptr(out), em)
        return ok, np.array(atten[:], dtype=np.float32), out[0], np.array(em[:], dtype=np.float32)


class RefOracle:
    """The reference's own CPU implementation (world->hit, color, getRay, textures...)."""

    def __init__(self, dof=False):
        so = REF_DOF_SO if dof else REF_SO
        if not os.path.exists(so):
            raise RuntimeError(f"{so} missing: run `make -C oracle` where /root/reference exists")
        L = C.CDLL(so)
        vp, u64 = C.c_void_p, C.c_uint64
        L.ref_last_error.restype = C.c_char_p
        L.ref_set_image_dir.argtypes = [C.c_char_p]
        L.ref_scene_build.restype = vp
        L.ref_scene_build.argtypes = [C.c_char_p, u64, C.c_int, C.c_int]
        L.ref_scene_free.argtypes = [vp]
        L.ref_scene_desc.restype = C.POINTER(abi.SceneDesc)
        L.ref_scene_desc.argtypes = [vp]
        L.ref_trace.restype = C.c_double
        L.ref_trace.argtypes = [vp, vp, u64, vp, u64, C.c_int]
        L.ref_camera_create.restype = vp
        L.ref_camera_create.argtypes = [C.c_float * 3, C.c_float * 3, C.c_float * 3] + [C.c_float] * 6
        L.ref_camera_free.argtypes = [vp]
        L.ref_camera_pod.argtypes = [vp, C.POINTER(abi.Camera)]
        L.ref_camera_pod_private.argtypes = [vp, C.POINTER(abi.Camera)]
        L.ref_camera_get_ray.argtypes = [vp, C.c_float, C.c_float, u64, vp]
        L.ref_render.restype = C.c_double
        L.ref_render.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, u64, C.c_int, vp, vp, vp,
                                 C.POINTER(u64), C.POINTER(u64)]
        L.ref_tonemap_bgra8.argtypes = [vp, C.c_int, C.c_int, vp]
        L.ref_color.argtypes = [vp, vp, u64, u64, vp]
        L.ref_aabb_hit.restype = C.c_int
        L.ref_aabb_hit.argtypes = [C.c_float * 3, C.c_float * 3, vp]
        L.ref_perlin_turbulence.argtypes = [vp, u64, vp]
        L.ref_perlin_noise.argtypes = [vp, u64, vp]
        L.ref_texture_value.restype = C.c_int
        L.ref_texture_value.argtypes = [vp, C.c_int, vp, vp, u64, vp]
        L.ref_sphere_uv.argtypes = [vp, u64, vp]
        L.ref_material_scatter.restype = C.c_int
        L.ref_material_scatter.argtypes = [vp, C.c_int, vp, vp, u64, C.c_float * 3, vp, C.c_float * 3]
        L.ref_image_texture_value.argtypes = [vp, C.c_int, C.c_int, vp, u64, vp]
        L.ref_reflect_refract.argtypes = [vp, vp, vp, u64, vp, vp, vp]
        L.ref_schlick.argtypes = [vp, vp, u64, vp]
        L.ref_decode_image.restype = C.POINTER(C.c_uint8)
        L.ref_decode_image.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ref_free_image.argtypes = [C.POINTER(C.c_uint8)]
        self.L = L
        L.ref_set_image_dir(REF_IMAGE_DIR.encode())

    def scene(self, name, seed=12345, n=0, wrap=True):
        h = self.L.ref_scene_build(name.encode(), seed, n, 1 if wrap else 0)
        if not h:
            raise RuntimeError("ref_scene_build: " + self.L.ref_last_error().decode())
        return RefScene(self, h, name)

    def camera(self, look_from, look_at, up, vfov, aspect, aperture, focus, t0=0.0, t1=1.0):
        f3 = C.c_float * 3
        return self.L.ref_camera_create(f3(*look_from), f3(*look_at), f3(*up), vfov, aspect, aperture, focus, t0, t1)

    def main_camera(self, W, H):
        """main.cpp:111-125"""
        return self.camera((0, 0, 0), (1, 0, 0), (0, 0, -1), 60.0, np.float32(W) / np.float32(H), 2.0, 1000.0, 0.0, 1.0)

    def camera_pod(self, cam, private=False):
        out = abi.Camera()
        (self.L.ref_camera_pod_private if private else self.L.ref_camera_pod)(cam, C.byref(out))
        return out

    def camera_ray(self, cam, s, t, seed=0):
        out = np.zeros(1, dtype=abi.ray_dtype)
        self.L.ref_camera_get_ray(cam, s, t, seed, _ptr(out))
        return out[0]

    def aabb_hit(self, bmin, bmax, ray):
        f3 = C.c_float * 3
        ray = np.ascontiguousarray(ray, dtype=abi.ray_dtype)
        return self.L.ref_aabb_hit(f3(*bmin), f3(*bmax), _ptr(ray))

    def perlin_turbulence(self, p):
        p = np.ascontiguousarray(p, dtype=np.float32)
        out = np.zeros(len(p), dtype=np.float32)
        self.L.ref_perlin_turbulence(_ptr(p), len(p), _ptr(out))
        return out

    def perlin_noise(self, p):
        p = np.ascontiguousarray(p, dtype=np.float32)
        out = np.zeros(len(p), dtype=np.float32)
        self.L.ref_perlin_noise(_ptr(p), len(p), _ptr(out))
        return out

    def sphere_uv(self, p):
        p = np.ascontiguousarray(p, dtype=np.float32)
        out = np.zeros((len(p), 2), dtype=np.float32)
        self.L.ref_sphere_uv(_ptr(p), len(p), _ptr(out))
        return out

    def image_texture_value(self, image, uv):
        """ImageTexture::value (texture.h:65-79) on caller-supplied RGB8 texels."""
        img = np.ascontiguousarray(image, dtype=np.uint8)
        uv = np.ascontiguousarray(uv, dtype=np.float32)
        out = np.zeros((len(uv), 3), dtype=np.float32)
        self.L.ref_image_texture_value(_ptr(img), img.shape[1], img.shape[0], _ptr(uv), len(uv), _ptr(out))
        return out

    def reflect_refract(self, v, n, ni):
        """material.h:11-26 -> (reflected, refracted, refract_ok)"""
        v = np.ascontiguousarray(v, dtype=np.float32); n = np.ascontiguousarray(n, dtype=np.float32)
        ni = np.ascontiguousarray(ni, dtype=np.float32)
        refl = np.zeros_like(v); refr = np.zeros_like(v); ok = np.zeros(len(v), dtype=np.int32)
        self.L.ref_reflect_refract(_ptr(v), _ptr(n), _ptr(ni), len(v), _ptr(refl), _ptr(refr), _ptr(ok))
        return refl, refr, ok

    def schlick(self, cosine, ri):
        cosine = np.ascontiguousarray(cosine, dtype=np.float32); ri = np.ascontiguousarray(ri, dtype=np.float32)
        out = np.zeros(len(cosine), dtype=np.float32)
        self.L.ref_schlick(_ptr(cosine), _ptr(ri), len(cosine), _ptr(out))
        return out

    def tonemap_bgra8(self, mean_rgb):
        mean_rgb = np.ascontiguousarray(mean_rgb, dtype=np.float32)
        H, W = mean_rgb.shape[:2]
        out = np.zeros((H, W, 4), dtype=np.uint8)
        self.L.ref_tonemap_bgra8(_ptr(mean_rgb), W, H, _ptr(out))
        return out

    def decode_image(self, path):
        nx, ny, nn = C.c_int(), C.c_int(), C.c_int()
        p = self.L.ref_decode_image(str(path).encode(), C.byref(nx), C.byref(ny), C.byref(nn))
        if not p:
            return None
        arr = np.ctypeslib.as_array(p, shape=(ny.value, nx.value, nn.value)).copy()
        self.L.ref_free_image(p)
        return arr


class PortOracle:
    """The plain-C restatement, operating on a flattened scene (abi.SceneDesc)."""

    def __init__(self):
        if not os.path.exists(PORT_SO):
            build()
        L = C.CDLL(PORT_SO)
        vp, u64 = C.c_void_p, C.c_uint64
        SD = C.POINTER(abi.SceneDesc)
        L.rtow_oracle_philox.argtypes = [C.c_uint32 * 4, C.c_uint32 * 2, C.c_uint32 * 4]
        L.rtow_oracle_aabb_hit.restype = C.c_int
        L.rtow_oracle_aabb_hit.argtypes = [C.c_float * 3, C.c_float * 3, C.c_float * 3, C.c_float * 3, C.c_float, C.c_float]
        L.rtow_oracle_trace.restype = C.c_int
        L.rtow_oracle_trace.argtypes = [SD, vp, u64, vp, u64, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.rtow_oracle_perlin_noise.restype = C.c_float
        L.rtow_oracle_perlin_noise.argtypes = [vp, C.c_float * 3]
        L.rtow_oracle_perlin_turbulence.restype = C.c_float
        L.rtow_oracle_perlin_turbulence.argtypes = [vp, C.c_float * 3]
        L.rtow_oracle_texture.argtypes = [SD, C.c_int, C.c_float, C.c_float, C.c_float * 3, C.c_float * 3]
        L.rtow_oracle_sphere_uv.argtypes = [C.c_float * 3, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.rtow_oracle_render.restype = C.c_int
        L.rtow_oracle_render.argtypes = [SD, C.POINTER(abi.Camera), C.POINTER(abi.RenderParams), vp, vp, C.POINTER(u64)]
        L.rtow_oracle_camera_ray.argtypes = [C.POINTER(abi.Camera), C.c_float, C.c_float, u64, vp]
        L.rtow_oracle_scatter_fixed.restype = C.c_int
        L.rtow_oracle_scatter_fixed.argtypes = [SD, vp, vp, u64, C.c_float * 3, vp, C.c_float * 3]
        f3 = C.c_float * 3
        L.rtow_oracle_reflect.argtypes = [f3, f3, f3]
        L.rtow_oracle_refract.restype = C.c_int
        L.rtow_oracle_refract.argtypes = [f3, f3, C.c_float, f3]
        L.rtow_oracle_schlick.restype = C.c_float
        L.rtow_oracle_schlick.argtypes = [C.c_float, C.c_float]
        L.rtow_oracle_resolve.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp]
        self.L = L

    def reflect_refract(self, v, n, ni):
        f3 = C.c_float * 3
        v = np.asarray(v, dtype=np.float32); n = np.asarray(n, dtype=np.float32)
        refl = np.zeros_like(v); refr = np.zeros_like(v); ok = np.zeros(len(v), dtype=np.int32)
        o = f3()
        for i in range(len(v)):
            self.L.rtow_oracle_reflect(f3(*v[i]), f3(*n[i]), o); refl[i] = o[:]
            ok[i] = self.L.rtow_oracle_refract(f3(*v[i]), f3(*n[i]), float(ni[i]), o); refr[i] = o[:]
        return refl, refr, ok

    def schlick(self, cosine, ri):
        return np.array([self.L.rtow_oracle_schlick(float(c), float(r)) for c, r in zip(cosine, ri)], dtype=np.float32)

    def resolve(self, sum_rgb, spp, want_bgra=True):
        sum_rgb = np.ascontiguousarray(sum_rgb, dtype=np.float32)
        H, W = sum_rgb.shape[:2]
        mean = np.zeros((H, W, 3), dtype=np.float32)
        bgra = np.zeros((H, W, 4), dtype=np.uint8)
        self.L.rtow_oracle_resolve(_ptr(sum_rgb), W, H, spp, _ptr(mean), _ptr(bgra))
        return mean, bgra

    def philox(self, ctr, key):
        out = (C.c_uint32 * 4)()
        self.L.rtow_oracle_philox((C.c_uint32 * 4)(*ctr), (C.c_uint32 * 2)(*key), out)
        return tuple(out)

    def aabb_hit(self, bmin, bmax, o, d, tmin, tmax):
        f3 = C.c_float * 3
        return self.L.rtow_oracle_aabb_hit(f3(*bmin), f3(*bmax), f3(*o), f3(*d), tmin, tmax)

    def trace(self, desc, rays, seed=0, counters=False):
        rays = np.ascontiguousarray(rays, dtype=abi.ray_dtype)
        hits = np.zeros(len(rays), dtype=abi.hit_dtype)
        nv, pt = C.c_double(), C.c_double()
        self.L.rtow_oracle_trace(C.byref(desc), _ptr(rays), len(rays), _ptr(hits), seed, C.byref(nv), C.byref(pt))
        return (hits, nv.value, pt.value) if counters else hits

    def perlin_noise(self, desc, p):
        f3 = C.c_float * 3
        return np.array([self.L.rtow_oracle_perlin_noise(desc.perlin, f3(*q)) for q in np.asarray(p, dtype=np.float32)],
                        dtype=np.float32)

    def perlin_turbulence(self, desc, p):
        f3 = C.c_float * 3
        return np.array([self.L.rtow_oracle_perlin_turbulence(desc.perlin, f3(*q)) for q in np.asarray(p, dtype=np.float32)],
                        dtype=np.float32)

    def texture_value(self, desc, tex_id, uv, p):
        f3 = C.c_float * 3
        out = np.zeros((len(p), 3), dtype=np.float32)
        rgb = f3()
        for i, (q, w) in enumerate(zip(np.asarray(p, dtype=np.float32), np.asarray(uv, dtype=np.float32))):
            self.L.rtow_oracle_texture(C.byref(desc), tex_id, float(w[0]), float(w[1]), f3(*q), rgb)
            out[i] = rgb[:]
        return out

    def sphere_uv(self, p):
        f3 = C.c_float * 3
        out = np.zeros((len(p), 2), dtype=np.float32)
        u, v = C.c_float(), C.c_float()
        for i, q in enumerate(np.asarray(p, dtype=np.float32)):
            self.L.rtow_oracle_sphere_uv(f3(*q), C.byref(u), C.byref(v))
            out[i] = (u.value, v.value)
        return out

    def render(self, desc, cam, params, want_sumsq=False):
        W, H = params.width, params.height
        acc = np.zeros((H, W, 3), dtype=np.float32)
        sq = np.zeros((H, W, 3), dtype=np.float64) if want_sumsq else None
        rays = C.c_uint64()
        self.L.rtow_oracle_render(C.byref(desc), C.byref(cam), C.byref(params), _ptr(acc),
                                  _ptr(sq) if want_sumsq else None, C.byref(rays))
        return dict(sum=acc, sumsq=sq, rays=rays.value)

    def camera_ray(self, cam, s, t, seed=0):
        out = np.zeros(1, dtype=abi.ray_dtype)
        self.L.rtow_oracle_camera_ray(C.byref(cam), s, t, seed, _ptr(out))
        return out[0]

    def scatter(self, desc, hit, ray, seed):
        atten = (C.c_float * 3)()
        em = (C.c_float * 3)()
        out = np.zeros(1, dtype=abi.ray_dtype)
        ray = np.ascontiguousarray(ray, dtype=abi.ray_dtype)
        hit = np.ascontiguousarray(hit, dtype=abi.hit_dtype)
        ok = self.L.rtow_oracle_scatter_fixed(C.byref(desc), _ptr(hit), _ptr(ray), seed, atten, _ptr(out), em)
        return ok, np.array(atten[:], dtype=np.float32), out[0], np.array(em[:], dtype=np.float32)


# ray-set generators live in the product package (bench.py uses them too); re-exported for the tests
coherent_rays = load_package().raysets.coherent_rays
incoherent_rays = load_package().raysets.incoherent_rays
camera_rays = load_package().raysets.camera_rays
