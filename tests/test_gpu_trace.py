"""-m gpu: closest hit through the C ABI (rtow_trace_rays) vs the oracle, bit for bit.
Gate (BASELINE.json north_star): primitive id and material id bit-exact; t, normal within 1e-5
relative -- here t, point and normal are required to be BIT-IDENTICAL; u,v within 1e-6 (atan2f/asinf)."""
import numpy as np
import pytest

from conftest import SCENES, bits, earth_image

pytestmark = pytest.mark.gpu


def _rays_for(ob, pkg, name, n=1 << 15):
    cam = pkg.main_camera(400, 225)
    ext = 100.0 if name == "bvh_stress" else 600.0
    sets = [ob.camera_rays(cam, 400, 225, n, seed=11), ob.incoherent_rays(n, seed=12, extent=ext)]
    if name == "bvh_stress":
        sets.append(ob.coherent_rays(128))
        sets.append(ob.coherent_rays(128, origin=(0.0, 0.0, -120.0)))
    return np.concatenate(sets)


def _compare(got, want, label):
    assert np.array_equal(got["prim_id"], want["prim_id"]), f"{label}: prim ids differ at {np.flatnonzero(got['prim_id'] != want['prim_id'])[:8]}"
    assert np.array_equal(got["mat_id"], want["mat_id"]), f"{label}: material ids differ"
    hit = want["prim_id"] >= 0
    for k in ("t", "p", "n"):
        assert np.array_equal(bits(got[k][hit]), bits(want[k][hit])), f"{label}: {k} not bit-identical"
    for k in ("u", "v"):
        assert np.allclose(got[k][hit], want[k][hit], rtol=0, atol=1e-6, equal_nan=True), f"{label}: {k} differs"  # asinf(|y|>1) is NaN on both sides


@pytest.mark.parametrize("name,kw", SCENES)
@pytest.mark.parametrize("mode", ["exact", "fast"])
def test_scene_hits_match_oracle(gpu, abi, ob, port, name, kw, mode):
    host = gpu.HostScene(name, seed=12345, image=earth_image(), **kw)
    dev = gpu.DeviceScene(host.desc)
    rays = _rays_for(ob, gpu, name)
    want = port.trace(host.desc, rays, seed=21)
    got, st = dev.trace(rays, traverse=abi.TRAVERSE_EXACT if mode == "exact" else abi.TRAVERSE_FAST, seed=21)
    assert st.rays == len(rays)
    _compare(got, want, f"{name}/{mode}")


@pytest.mark.parametrize("n", [1, 2, 3, 1000, 20000])
@pytest.mark.parametrize("mode", ["exact", "fast"])
def test_bvh_stress_hits_match_oracle(gpu, abi, ob, port, n, mode):
    host = gpu.HostScene("bvh_stress", seed=777, n=n)
    dev = gpu.DeviceScene(host.desc)
    rays = _rays_for(ob, gpu, "bvh_stress", n=1 << 14)
    want = port.trace(host.desc, rays, seed=1)
    got, _ = dev.trace(rays, traverse=abi.TRAVERSE_EXACT if mode == "exact" else abi.TRAVERSE_FAST, seed=1)
    _compare(got, want, f"bvh_stress[{n}]/{mode}")


@pytest.mark.parametrize("n_spheres,log2_rays", [(200000, 18), (1000000, 21)])
def test_fast_equals_exact_on_large_bvh(gpu, abi, ob, n_spheres, log2_rays):
    """Size-independent property at a size the CPU oracle cannot sweep: the culled traversal of the
    rebuilt BVH returns exactly what the un-culled walk of the reference tree returns."""
    host = gpu.HostScene("bvh_stress", seed=4242, n=n_spheres)
    dev = gpu.DeviceScene(host.desc)
    rays = np.concatenate([ob.coherent_rays(512), ob.coherent_rays(512, origin=(0.0, 0.0, -120.0)),
                           ob.incoherent_rays(1 << log2_rays, seed=5, extent=100.0), ob.incoherent_rays(1 << (log2_rays - 2), seed=6, extent=300.0)])
    a, _ = dev.trace(rays, traverse=abi.TRAVERSE_EXACT)
    b, _ = dev.trace(rays, traverse=abi.TRAVERSE_FAST)
    assert (a["prim_id"] >= 0).mean() > 0.2
    assert a.tobytes() == b.tobytes()
    nv_e, pt_e = dev.trace_counters(rays[:1 << 16], traverse=abi.TRAVERSE_EXACT)
    nv_f, pt_f = dev.trace_counters(rays[:1 << 16], traverse=abi.TRAVERSE_FAST)
    assert nv_f < nv_e and pt_f < pt_e


def test_edge_rays(gpu, abi, ob, port):
    """Empty batch, zero / axis-aligned directions, negative tmin, tmax cut-offs, NaN-free misses."""
    host = gpu.HostScene("bvh_stress", seed=9, n=5000)
    dev = gpu.DeviceScene(host.desc)
    got, st = dev.trace(np.zeros(0, dtype=abi.ray_dtype))
    assert len(got) == 0 and st.rays == 0
    rays = ob.incoherent_rays(4096, seed=3, extent=100.0)
    rays["d"][0::4, 0] = 0.0                      # zero component
    rays["d"][1::4] = (0.0, 0.0, 1.0)             # axis aligned
    rays["tmin"][2::4] = -50.0                    # negative tmin
    rays["tmax"][3::4] = 15.0                     # finite tmax
    want = port.trace(host.desc, rays)
    for mode in (abi.TRAVERSE_EXACT, abi.TRAVERSE_FAST):
        got, _ = dev.trace(rays, traverse=mode)
        _compare(got, want, f"edge/{mode}")


def test_counters_match_oracle(gpu, abi, ob, port):
    host = gpu.HostScene("random_ned", seed=12345, image=earth_image())
    dev = gpu.DeviceScene(host.desc)
    rays = ob.camera_rays(gpu.main_camera(400, 225), 400, 225, 8192, seed=2)
    _, nv, pt = port.trace(host.desc, rays, counters=True)
    g_nv, g_pt = dev.trace_counters(rays, traverse=abi.TRAVERSE_EXACT)
    # node (slab) tests are the reference's count exactly; an n==1 BvhNode tests its leaf twice in the reference
    # (bvhNode.h:38-40,70-71) and the oracle counts both, the kernel evaluates a non-medium leaf once (same answer)
    assert abs(g_nv - nv) < 1e-9 and 0 < g_pt <= pt + 1e-9


def test_host_batches_are_pipelined_in_chunks_without_changing_hits(gpu, abi, ob):
    """rtow_trace_rays sends a large host batch through in chunks of 2^20 rays on two streams (copies under kernels).  A medium's
    free-flight draw is keyed by the ray's index in the WHOLE batch, so the chunked call must return exactly what one
    unchunked launch over the same rays returns (rtow_trace_rays_device), media hits included."""
    import torch
    host = gpu.HostScene("cornell_smoke_ned", seed=12345, wrap=True)
    dev = gpu.DeviceScene(host.desc)
    n = (1 << 21) + 12345                        # three chunks, the last one ragged
    rays = np.concatenate([ob.camera_rays(gpu.main_camera(640, 360), 640, 360, 1 << 20, seed=3), ob.incoherent_rays(n - (1 << 20), seed=4, extent=450.0)])
    got, st = dev.trace(rays, seed=77)
    assert st.rays == n
    d_rays = torch.from_numpy(rays.view(np.uint8).reshape(n, -1).copy()).cuda()
    d_hits = torch.empty(n, 48, dtype=torch.uint8, device="cuda:0")
    dev.trace_device(d_rays.data_ptr(), n, d_hits.data_ptr(), abi.TRAVERSE_AUTO, 77, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    want = d_hits.cpu().numpy().view(abi.hit_dtype).reshape(-1)
    mats = host.arrays()["materials"]
    assert (mats["type"][np.maximum(want["mat_id"], 0)][want["mat_id"] >= 0] == abi.MAT_ISOTROPIC).sum() > 1000, "the batch must exercise media"
    assert got.tobytes() == want.tobytes()
