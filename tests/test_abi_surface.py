"""not-gpu: the C-ABI library loads, exports every entry point include/*.h declares, its PODs have the sizes the
Python mirrors assume, and -- on a box without a CUDA device -- compute entry points fail loudly (no CPU path)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_entry_points():
    names = []
    for hdr in ("rtow_b200.h", "rtow_b200_scenes.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"\b(rtow_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_every_declared_symbol_is_exported(pkg):
    names = declared_entry_points()
    assert len(names) >= 20 and "rtow_render" in names and "rtow_trace_rays" in names
    raw = C.CDLL(pkg.lib_path())
    missing = [n for n in names if not hasattr(raw, n)]
    assert not missing, f"declared in include/*.h but not exported: {missing}"
    pkg.lib()   # binds argtypes for every symbol; raises if one is absent


def test_pod_sizes_match_the_header(pkg, abi):
    text = open(os.path.join(ROOT, "include", "rtow_b200.h")).read()
    assert "/* 36 B */" in text and "/* 48 B = 3 x float4 */" in text and "/* 80 B */" in text
    assert abi.ray_dtype.itemsize == 36 and abi.hit_dtype.itemsize == 48 and abi.prim_dtype.itemsize == 80
    assert abi.node_dtype.itemsize == 32 and C.sizeof(abi.Camera) == 100 and C.sizeof(abi.Stats) == 32
    assert C.sizeof(abi.RenderParams) == 72 and abi.perlin_dtype.itemsize == 256 * 12 + 3 * 1024
    assert b"sm_100a" in pkg.lib().rtow_build_info()


def test_host_entry_points_work_without_a_device(pkg, abi, tmp_path):
    """Scene building, flattening, camera set-up and the image writers are host code."""
    hs = pkg.HostScene("cornell_ned", seed=1)
    a = hs.arrays()
    assert a["header"]["top_kind"] == abi.TOP_LIST and len(a["prims"]) > 12 and len(a["nodes"]) == 0
    hb = pkg.HostScene("bvh_stress", seed=1, n=100)
    b = hb.arrays()
    assert b["header"]["top_kind"] == abi.TOP_BVH and len(b["objects"]) == 100 and len(b["nodes"]) >= 50
    with pytest.raises(pkg.RtowError):
        pkg.HostScene("no_such_scene")
    with pytest.raises(pkg.RtowError):
        pkg.HostScene("bvh_stress", n=0)
    img = np.linspace(0, 1.2, 6 * 4 * 3, dtype=np.float32).reshape(4, 6, 3)
    pkg.write_ppm(tmp_path / "a.ppm", img)
    pkg.write_pfm(tmp_path / "a.pfm", img)
    ppm = open(tmp_path / "a.ppm", "rb").read()
    assert ppm.startswith(b"P6\n6 4\n255\n") and len(ppm) == 11 + 72
    want = np.minimum((255.99 * np.sqrt(img).astype(np.float64)).astype(np.int64), 255).astype(np.uint8)   # main.cpp:956-970
    assert ppm[11:] == want.tobytes()
    pfm = open(tmp_path / "a.pfm", "rb").read()
    assert pfm.startswith(b"PF\n6 4\n-1.0\n") and np.array_equal(np.frombuffer(pfm[12:], dtype="<f4").reshape(4, 6, 3)[::-1], img)


def _read_exr(path):
    """Minimal reader for single-part, uncompressed, FLOAT scanline OpenEXR (the subset rtow_write_exr emits)."""
    import struct
    b = open(path, "rb").read()
    assert struct.unpack_from("<ii", b, 0) == (20000630, 2)
    pos, attrs = 8, {}
    while b[pos] != 0:
        e = b.index(b"\0", pos); name = b[pos:e].decode(); pos = e + 1
        e = b.index(b"\0", pos); typ = b[pos:e].decode(); pos = e + 1
        size, = struct.unpack_from("<i", b, pos); pos += 4
        attrs[name] = (typ, b[pos:pos + size]); pos += size
    pos += 1
    chans, q, raw = [], 0, attrs["channels"][1]
    while raw[q] != 0:
        e = raw.index(b"\0", q); nm = raw[q:e].decode(); q = e + 1
        ptype, plin, xs, ys = struct.unpack_from("<iiii", raw, q); q += 16
        assert ptype == 2 and xs == ys == 1
        chans.append(nm)
    x0, y0, x1, y1 = struct.unpack("<iiii", attrs["dataWindow"][1])
    W, H = x1 - x0 + 1, y1 - y0 + 1
    assert attrs["compression"][1] == b"\0" and attrs["lineOrder"][1] == b"\0" and attrs["displayWindow"][1] == attrs["dataWindow"][1]
    offs = struct.unpack_from("<%dQ" % H, b, pos)
    img = np.zeros((H, W, 3), dtype=np.float32)
    for row, off in enumerate(offs):
        y, size = struct.unpack_from("<ii", b, off)
        assert y == row and size == W * 4 * len(chans)
        data = np.frombuffer(b, dtype="<f4", count=W * len(chans), offset=off + 8).reshape(len(chans), W)
        for c, nm in enumerate(chans):
            img[row, :, "RGB".index(nm)] = data[c]
    assert offs[-1] + 8 + W * 4 * len(chans) == len(b)
    return chans, img


def test_exr_writer_and_oriented_camera(pkg, abi, tmp_path):
    """hand-rolled OpenEXR output; per-frame camera update (camera.h:166-229) as a fresh POD."""
    img = np.random.default_rng(3).uniform(0, 9, (5, 7, 3)).astype(np.float32)
    pkg.write_exr(tmp_path / "a.exr", img)
    chans, back = _read_exr(tmp_path / "a.exr")
    assert chans == ["B", "G", "R"] and np.array_equal(back, img)
    with pytest.raises(pkg.RtowError):
        pkg.write_exr(tmp_path / "no_such_dir" / "a.exr", img)
    # the constructor's hard-coded NED frame (camera.h:76-81) passed explicitly gives main()'s camera back, byte for byte
    ned = np.array([[-1, 0, 0, 0], [0, 1, 0, 0], [0, 0, -1, 0], [0, 0, 0, 1]], dtype=np.float32)
    assert bytes(pkg.oriented_camera((0, 0, 0), ned, 60.0, 16 / 9)) == bytes(pkg.main_camera(1920, 1080))
    # a yaw of the frame and a moved eye change origin and screen vectors but not their lengths
    c, s_ = np.cos(0.3), np.sin(0.3)
    yaw = np.array([[-c, s_, 0, 0], [s_, c, 0, 0], [0, 0, -1, 0], [0, 0, 0, 1]], dtype=np.float32)
    a, b = pkg.main_camera(1920, 1080), pkg.oriented_camera((5, -2, 1), yaw, 60.0, 16 / 9)
    fa, fb = np.frombuffer(bytes(a), dtype=np.float32), np.frombuffer(bytes(b), dtype=np.float32)
    assert not np.array_equal(fa, fb)
    assert np.allclose(fb[0:3], (5, -2, 1))
    assert np.isclose(np.linalg.norm(fa[6:9]), np.linalg.norm(fb[6:9]), rtol=1e-6) and np.isclose(np.linalg.norm(fa[9:12]), np.linalg.norm(fb[9:12]), rtol=1e-6)


def test_cpp_convenience_layer_compiles(pkg, tmp_path):
    """include/rtow/render.h (the C++ convenience over the C ABI): the example builds and links against the
    library, and -- where the reference sources exist -- the same header instantiates on the reference's OWN classes, which
    is the binding INTEGRATION.md tells a maintainer to add."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.dirname(pkg.lib_path())
    exe = str(tmp_path / "render_api")
    r = subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", os.path.join(root, "examples", "render_api.cpp"), "-I" + os.path.join(root, "include"),
                        "-L" + libdir, "-lrtow_b200", "-Wl,-rpath," + libdir, "-Wl,-rpath-link,/usr/local/cuda/lib64", "-o", exe],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    if pkg.lib().rtow_device_count() == 0:
        run = subprocess.run([exe, "32", "18", "2"], capture_output=True, text=True)
        assert run.returncode == 1 and "no CPU path" in run.stderr
    ref = "/root/reference/rayTracingOneWeekend"
    if not os.path.isdir(ref):
        return
    tu = tmp_path / "ref_binding.cpp"
    tu.write_text("""
#include <cfloat>
#include "scenes.h"
#include "rtow/render.h"
struct RefTraits {
    using vec3 = ::vec3; using Hitable = ::Hitable; using Material = ::Material; using Texture = ::Texture;
    using FlipNormals = ::FlipNormals; using Translate = ::Translate; using RotateY = ::RotateY;
    using Sphere = ::Sphere; using MovingSphere = ::MovingSphere;
    using XYRectangle = ::XYRectangle; using XZRectangle = ::XZRectangle; using YZRectangle = ::YZRectangle;
    using Box = ::Box; using ConstantMedium = ::ConstantMedium; using HitableList = ::HitableList; using BvhNode = ::BvhNode;
    using Lambertian = ::Lambertian; using Metal = ::Metal; using Dielectric = ::Dielectric;
    using DiffuseLight = ::DiffuseLight; using Isotropic = ::Isotropic;
    using ConstantTexture = ::ConstantTexture; using CheckerTexture = ::CheckerTexture;
    using NoiseTexture = ::NoiseTexture; using ImageTexture = ::ImageTexture;
    using Perlin = ::Perlin; using Camera = ::Camera;
    static double unif01() { return unifRand(randomNumberGenerator); }
};
rtow_stats frame(Hitable *world, Camera *cam, rtow::Image &out) {
    rtow::DeviceWorld<RefTraits> gpu(world);
    return gpu.render(*cam, rtow::default_params(800, 800, 64), out);
}
rtow_stats frame_on_all_gpus(Hitable *world, Camera *cam, rtow::Image &out) {
    rtow::MultiDeviceWorld<RefTraits> gpus(world);      // main.cpp:182-195,936 -> one host thread per GPU behind the C ABI
    return gpus.render(*cam, rtow::default_params(800, 800, 64), out);
}
""")
    r = subprocess.run(["g++", "-std=c++17", "-w", "-fsyntax-only", str(tu), "-I" + ref, "-I" + os.path.join(root, "include")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr


def test_compute_fails_loudly_without_a_device(pkg, abi):
    if pkg.lib().rtow_device_count() > 0:
        pytest.skip("a CUDA device is visible")
    hs = pkg.HostScene("cornell_ned", seed=1)
    with pytest.raises(pkg.RtowError) as e:
        pkg.DeviceScene(hs.desc)
    assert e.value.code == abi.ERR_NO_DEVICE and "no CPU path" in str(e.value)


def test_product_never_touches_the_oracle():
    """The oracle is test infrastructure: nothing under the package or include/ may reference it."""
    bad = []
    for base in ("ray-tracing-one-weekend_b200", "include"):
        for dp, _, fns in os.walk(os.path.join(ROOT, base)):
            for fn in fns:
                if fn.endswith((".py", ".h", ".cuh", ".cu", ".cpp")):
                    t = open(os.path.join(dp, fn), errors="replace").read()
                    if re.search(r'^\s*(from|import)\s+oracle|#include\s+"[^"]*oracle|librtow_oracle|libref_oracle|rtow_oracle_\w+\s*\(', t, flags=re.M):
                        bad.append(os.path.join(dp, fn))
    assert not bad, bad


def test_cli_fails_loudly_without_a_device(pkg):
    import subprocess
    cli = os.path.join(ROOT, "ray-tracing-one-weekend_b200", "csrc", "rtow_cli")
    assert os.path.exists(cli), "rtow_cli is built by __graft_entry__.build()"
    if pkg.lib().rtow_device_count() > 0:
        pytest.skip("a CUDA device is visible")
    out = subprocess.run([cli, "--scene", "cornell_ned", "--width", "8", "--height", "8", "--spp", "1"], capture_output=True, text=True)
    assert out.returncode == 3 and "no CPU path" in out.stderr


def test_scene_create_rejects_a_dag_and_out_of_order_leaves(pkg, abi):
    """rtow_scene_create validates before it touches a device: a BVH given as a DAG (a node or object under two parents) or with
    leaves not numbered along the in-order (the tie rule of bvhNode.h:74-77 is implemented as 'larger object id wins') is
    RTOW_ERR_INVALID_ARG.  A well-formed scene passes validation and fails later only for want of a device (on CPU boxes)."""
    host = pkg.HostScene("bvh_stress", seed=5, n=9)
    base = host.arrays()["nodes"].copy()
    assert len(base) >= 4

    def create(nodes):
        d = abi.SceneDesc.from_buffer_copy(bytes(host.desc))      # a private copy of the POD (pointers to the scene's arrays included)
        keep = np.ascontiguousarray(nodes)
        d.nodes = keep.ctypes.data
        h = C.c_void_p()
        rc = pkg.lib().rtow_scene_create(C.byref(d), 0, C.byref(h))
        msg = pkg.lib().rtow_last_error().decode()
        if h:
            pkg.lib().rtow_scene_destroy(h)
        return rc, msg

    rc, msg = create(base)
    assert rc in (abi.OK, abi.ERR_NO_DEVICE), msg
    internal = [i for i in range(len(base)) if base["left"][i] >= 0 and base["right"][i] >= 0]
    dag = base.copy()
    i = internal[0]
    dag["right"][i] = dag["left"][i]                    # both children the same subtree
    rc, msg = create(dag)
    assert rc == abi.ERR_INVALID_ARG, msg
    dag = base.copy()
    leaf_nodes = [i for i in range(len(base)) if base["left"][i] < 0 and base["right"][i] < 0 and base["left"][i] != base["right"][i]]
    a, b = leaf_nodes[0], leaf_nodes[-1]
    dag["right"][b] = dag["left"][a]                    # one object under two parents
    rc, msg = create(dag)
    assert rc == abi.ERR_INVALID_ARG and ("more than one parent" in msg or "in-order" in msg), msg
    swapped = base.copy()
    swapped["left"][a], swapped["right"][a] = base["right"][a], base["left"][a]      # leaves out of in-order
    rc, msg = create(swapped)
    assert rc == abi.ERR_INVALID_ARG and "in-order" in msg, msg
