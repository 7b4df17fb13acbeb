"""ctypes / numpy mirrors of the POD structs in include/rtow_b200.h.

Pure layout code -- no computation.  Shared by the product binding (this package) and by the
test-only oracle bindings under oracle/.  Sizes are asserted against the C side at load time
(`rtow_abi_sizes`).
"""
import ctypes as C

import numpy as np

ABI_VERSION = 1

# status codes
OK, ERR_INVALID_ARG, ERR_UNSUPPORTED, ERR_NO_DEVICE, ERR_CUDA, ERR_OOM, ERR_IO = 0, -1, -2, -3, -4, -5, -6

XF_TRANSLATE, XF_ROTATE_Y, XF_FLIP = 0, 1, 2
PRIM_SPHERE, PRIM_MOVING_SPHERE, PRIM_RECT_XY, PRIM_RECT_XZ, PRIM_RECT_YZ, PRIM_MEDIUM = range(6)
MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_DIFFUSE_LIGHT, MAT_ISOTROPIC = range(5)
TEX_CONSTANT, TEX_CHECKER, TEX_NOISE, TEX_IMAGE = range(4)
TOP_LIST, TOP_BVH = 0, 1
KERNEL_AUTO, KERNEL_MEGA, KERNEL_WAVEFRONT = 0, 1, 2
TRAVERSE_AUTO, TRAVERSE_EXACT, TRAVERSE_FAST = 0, 1, 2

FLT_MAX = float(np.finfo(np.float32).max)

ray_dtype = np.dtype(
    [("o", "<f4", 3), ("d", "<f4", 3), ("time", "<f4"), ("tmin", "<f4"), ("tmax", "<f4")], align=False
)
hit_dtype = np.dtype(
    [("t", "<f4"), ("prim_id", "<i4"), ("mat_id", "<i4"), ("n", "<f4", 3), ("p", "<f4", 3), ("u", "<f4"), ("v", "<f4"), ("reserved", "<i4")],
    align=False,
)
xform_dtype = np.dtype([("type", "<i4"), ("f", "<f4", 3)])
prim_dtype = np.dtype(
    [("type", "<i4"), ("mat_id", "<i4"), ("xform_begin", "<i4"), ("xform_count", "<i4"), ("f", "<f4", 12),
     ("child_begin", "<i4"), ("child_count", "<i4"), ("object_id", "<i4"), ("reserved", "<i4")]
)
object_dtype = np.dtype([("prim_begin", "<i4"), ("prim_count", "<i4")])
node_dtype = np.dtype([("bmin", "<f4", 3), ("left", "<i4"), ("bmax", "<f4", 3), ("right", "<i4")])
material_dtype = np.dtype([("type", "<i4"), ("tex_id", "<i4"), ("f", "<f4", 4)])
texture_dtype = np.dtype([("type", "<i4"), ("a", "<i4"), ("b", "<i4"), ("f", "<f4", 3)])
perlin_dtype = np.dtype([("rand_vec", "<f4", (256, 3)), ("perm_x", "<i4", 256), ("perm_y", "<i4", 256), ("perm_z", "<i4", 256)])

assert ray_dtype.itemsize == 36 and hit_dtype.itemsize == 48 and prim_dtype.itemsize == 80
assert node_dtype.itemsize == 32 and xform_dtype.itemsize == 16


class Image(C.Structure):
    _fields_ = [("data", C.POINTER(C.c_uint8)), ("nx", C.c_int32), ("ny", C.c_int32)]


class SceneDesc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32),
        ("top_kind", C.c_int32),
        ("root_xform_begin", C.c_int32),
        ("root_xform_count", C.c_int32),
        ("root_node", C.c_int32),
        ("n_xforms", C.c_int32),
        ("n_prims", C.c_int32),
        ("n_objects", C.c_int32),
        ("n_nodes", C.c_int32),
        ("n_materials", C.c_int32),
        ("n_textures", C.c_int32),
        ("n_images", C.c_int32),
        ("xforms", C.c_void_p),
        ("prims", C.c_void_p),
        ("objects", C.c_void_p),
        ("nodes", C.c_void_p),
        ("materials", C.c_void_p),
        ("textures", C.c_void_p),
        ("images", C.POINTER(Image)),
        ("perlin", C.c_void_p),
    ]


class Camera(C.Structure):
    _fields_ = [
        ("origin", C.c_float * 3),
        ("lower_left", C.c_float * 3),
        ("horizontal", C.c_float * 3),
        ("vertical", C.c_float * 3),
        ("u", C.c_float * 3),
        ("v", C.c_float * 3),
        ("w", C.c_float * 3),
        ("lens_radius", C.c_float),
        ("time0", C.c_float),
        ("time1", C.c_float),
        ("dof_enabled", C.c_int32),
    ]

    def as_tuple(self):
        return tuple(
            tuple(getattr(self, k)) if hasattr(getattr(self, k), "__len__") else getattr(self, k)
            for k, _ in self._fields_
        )


class RenderParams(C.Structure):
    _fields_ = [
        ("width", C.c_int32),
        ("height", C.c_int32),
        ("spp", C.c_int32),
        ("sample_begin", C.c_int32),
        ("sample_count", C.c_int32),
        ("max_depth", C.c_int32),
        ("seed", C.c_uint64),
        ("global_illum_en", C.c_int32),
        ("sky_gain", C.c_float),
        ("ambient_gain", C.c_float),
        ("kernel", C.c_int32),
        ("traverse", C.c_int32),
        ("row_begin", C.c_int32),
        ("row_count", C.c_int32),
        ("reserved", C.c_int32 * 2),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("paths", C.c_uint64),
        ("rays", C.c_uint64),
        ("kernel_ms", C.c_float),
        ("total_ms", C.c_float),
        ("launches", C.c_uint32),
        ("reserved", C.c_uint32),
    ]


class RenderProfile(C.Structure):
    _fields_ = [
        ("kernel", C.c_int32),
        ("iterations", C.c_uint32),
        ("pool_slots", C.c_uint32),
        ("sampled_iterations", C.c_uint32),
        ("extend_ms_sum", C.c_float),
        ("shade_ms_sum", C.c_float),
        ("tail_iterations", C.c_uint32),
        ("polls", C.c_uint32),
        ("tail_ms", C.c_float),
        ("finish_launches", C.c_uint32),
    ]


class WorkCounters(C.Structure):
    _fields_ = [
        ("paths", C.c_uint64),
        ("rays", C.c_uint64),
        ("node_visits_per_ray", C.c_double),
        ("prim_tests_per_ray", C.c_double),
        ("noise_evals_per_ray", C.c_double),
        ("checker_evals_per_ray", C.c_double),
        ("image_evals_per_ray", C.c_double),
        ("scatters_per_ray", C.c_double),
        ("traverse", C.c_int32),
        ("reserved", C.c_int32),
    ]


def default_render_params(width, height, spp, seed=1, max_depth=50, kernel=KERNEL_AUTO, traverse=TRAVERSE_AUTO,
                          sample_begin=0, sample_count=0, row_begin=0, row_count=0):
    """defines.h defaults: GLOBAL_ILLUM_EN 1, SKY_ILLUM_GAIN 1.0, GLOBAL_ILLUM_GAIN 0.3, DEPTH_RECURSION 50."""
    p = RenderParams()
    p.width, p.height, p.spp = width, height, spp
    p.sample_begin, p.sample_count = sample_begin, sample_count
    p.max_depth = max_depth
    p.seed = seed
    p.global_illum_en = 1
    p.sky_gain = 1.0
    p.ambient_gain = 0.3
    p.kernel = kernel
    p.traverse = traverse
    p.row_begin, p.row_count = row_begin, row_count
    return p


def _view(ptr, count, dtype):
    if not ptr or count <= 0:
        return np.zeros(0, dtype=dtype)
    buf = (C.c_uint8 * (count * dtype.itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype, count=count)


def desc_arrays(desc):
    """numpy views (no copy) of every array a SceneDesc points to; valid while its owner lives."""
    out = {
        "xforms": _view(desc.xforms, desc.n_xforms, xform_dtype),
        "prims": _view(desc.prims, desc.n_prims, prim_dtype),
        "objects": _view(desc.objects, desc.n_objects, object_dtype),
        "nodes": _view(desc.nodes, desc.n_nodes, node_dtype),
        "materials": _view(desc.materials, desc.n_materials, material_dtype),
        "textures": _view(desc.textures, desc.n_textures, texture_dtype),
        "perlin": _view(desc.perlin, 1 if desc.perlin else 0, perlin_dtype),
    }
    images = []
    for i in range(desc.n_images):
        im = desc.images[i]
        n = im.nx * im.ny * 3
        addr = C.cast(im.data, C.c_void_p).value
        images.append(_view(addr, n, np.dtype("u1")).reshape(im.ny, im.nx, 3))
    out["images"] = images
    out["header"] = dict(top_kind=desc.top_kind, root_xform_begin=desc.root_xform_begin,
                         root_xform_count=desc.root_xform_count, root_node=desc.root_node)
    return out


class OwnedDesc:
    """A SceneDesc whose arrays are numpy copies held by this object (picklable via arrays())."""

    def __init__(self, arrays):
        self.a = {k: (np.ascontiguousarray(v) if isinstance(v, np.ndarray) else v) for k, v in arrays.items()}
        self.a["images"] = [np.ascontiguousarray(im) for im in arrays["images"]]
        d = SceneDesc()
        h = self.a["header"]
        d.abi_version = ABI_VERSION
        d.top_kind, d.root_xform_begin = h["top_kind"], h["root_xform_begin"]
        d.root_xform_count, d.root_node = h["root_xform_count"], h["root_node"]
        for name, key in (("xforms", "xforms"), ("prims", "prims"), ("objects", "objects"), ("nodes", "nodes"),
                          ("materials", "materials"), ("textures", "textures")):
            arr = self.a[key]
            setattr(d, "n_" + name, len(arr))
            setattr(d, name, arr.ctypes.data if len(arr) else None)
        self._imgs = (Image * max(1, len(self.a["images"])))()
        for i, im in enumerate(self.a["images"]):
            self._imgs[i].data = im.ctypes.data_as(C.POINTER(C.c_uint8))
            self._imgs[i].ny, self._imgs[i].nx = im.shape[0], im.shape[1]
        d.n_images = len(self.a["images"])
        d.images = C.cast(self._imgs, C.POINTER(Image))
        d.perlin = self.a["perlin"].ctypes.data if len(self.a["perlin"]) else None
        self.desc = d

    @classmethod
    def copy_of(cls, desc):
        arrs = desc_arrays(desc)
        return cls({k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in arrs.items() if k != "images"}
                   | {"images": [im.copy() for im in arrs["images"]]})
