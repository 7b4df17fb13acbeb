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
