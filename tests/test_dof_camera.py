"""Depth-of-field branch of Camera::getRay (camera.h:125-126: rd = lensRadius * randomInUnitDisk() + randomInUnitSphere(),
mathUtilities.h:11-25).  The reference ships with CAMERA_DOF_EN 0 (defines.h:26); oracle/_ref/libref_oracle_dof.so is the
same bridge compiled with the switch on (oracle/Makefile), so the branch has a live reference to stand against.

The reference draws from mt19937_64, the oracle and the kernels from Philox: ray origins are compared as distributions
(first and second moments of the lens offset), the chief-ray identity o + d = lowerLeft + s*H + t*V exactly-ish, and images
through the variance-normalised statistic the other image tests use."""
import numpy as np
import pytest


@pytest.fixture(scope="module")
def ref_dof(ob):
    if not ob.have_ref_dof():
        pytest.skip("oracle/_ref/libref_oracle_dof.so not built (needs /root/reference)")
    return ob.RefOracle(dof=True)


def _dof_camera(pkg, W, H):
    cam = pkg.main_camera(W, H)
    cam.dof_enabled = 1
    return cam


def test_dof_camera_pod_and_ray_distribution_match_the_reference(pkg, ob, ref_dof, port):
    W, H = 400, 225
    rc = ref_dof.main_camera(W, H)
    pod = ref_dof.camera_pod(rc)
    assert pod.dof_enabled == 1
    mine = _dof_camera(pkg, W, H)
    assert bytes(mine) == bytes(pod)
    n = 20000
    s, t = 0.37, 0.61
    a = np.stack([ref_dof.camera_ray(rc, s, t, seed=1000 + i) for i in range(n)])
    b = np.stack([port.camera_ray(mine, s, t, seed=1000 + i) for i in range(n)])
    origin = np.array(list(pod.origin), dtype=np.float64)
    for rays in (a, b):
        # the chief ray: origin + direction is the same point of the focus plane for every lens sample
        tgt = rays["o"].astype(np.float64) + rays["d"].astype(np.float64)
        assert np.abs(tgt - tgt[0]).max() <= 1e-3 * max(1.0, np.abs(tgt[0]).max())
    assert np.abs((a["o"][0].astype(np.float64) + a["d"][0]) - (b["o"][0].astype(np.float64) + b["d"][0])).max() <= 1e-3 * 2000
    oa, obb = a["o"].astype(np.float64) - origin, b["o"].astype(np.float64) - origin
    # offset = u * rd.x + v * rd.y with rd.x, rd.y ~ lensRadius * disk + sphere: zero mean, variance lr^2 / 4 + 1 / 5 per axis
    var_expected = float(pod.lens_radius) ** 2 / 4.0 + 0.2
    for off in (oa, obb):
        moved = off.std(axis=0) > 1e-6
        assert moved.sum() == 2                                   # the lens offset lives in the (u, v) plane
        assert np.abs(off[:, moved].mean(axis=0)).max() < 5 * np.sqrt(var_expected / n)
        assert np.allclose(off[:, moved].var(axis=0), var_expected, rtol=0.06)
    assert np.allclose(np.cov(oa.T), np.cov(obb.T), atol=0.06 * var_expected)
    # fourth moment too (a disk + a ball is not a Gaussian: wrong shapes with the right variance would pass otherwise)
    ka, kb = (oa ** 4).mean(axis=0), (obb ** 4).mean(axis=0)
    assert np.allclose(ka, kb, rtol=0.12, atol=1e-9)
    ta, tb = a["time"], b["time"]
    assert 0 <= ta.min() and ta.max() < 1 and 0 <= tb.min() and tb.max() < 1
    assert abs(ta.mean() - 0.5) < 0.02 and abs(tb.mean() - 0.5) < 0.02


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", ["mega", "wavefront"])
def test_dof_render_matches_oracle_and_reference(gpu, abi, ob, port, ref_dof, kernel):
    """GPU vs the port oracle: same Philox streams, so per-pixel sums agree to rounding and ray counts exactly.  GPU vs the
    reference built with CAMERA_DOF_EN 1: the same image up to Monte-Carlo noise."""
    from conftest import earth_image
    from test_oracle_golden import z2_statistic
    W, H, spp = 64, 36, 8
    k = abi.KERNEL_MEGA if kernel == "mega" else abi.KERNEL_WAVEFRONT
    host = gpu.HostScene("random_ned", seed=12345, wrap=True, image=earth_image())
    dev = gpu.DeviceScene(host.desc)
    cam = _dof_camera(gpu, W, H)
    p = abi.default_render_params(W, H, spp, seed=5, kernel=k)
    img, st = dev.render(cam, p)
    want = port.render(host.desc, cam, p)
    assert st.rays == want["rays"]
    ref = want["sum"] / np.float32(spp)
    assert np.abs(img - ref).max() <= 2e-4 * max(1.0, float(np.abs(ref).max()))
    plain, _ = dev.render(gpu.main_camera(W, H), p)
    assert not np.array_equal(plain, img), "dof_enabled changed nothing"

    # statistical leg against the reference's own DOF build, on the reference's own scene (its decoded JPEG texture)
    W, H, spp_gpu, spp_ref = 48, 27, 2048, 1024
    host = gpu.HostScene("random_ned", seed=12345, wrap=True)
    dev = gpu.DeviceScene(host.desc)
    cam = _dof_camera(gpu, W, H)
    img, st = dev.render(cam, abi.default_render_params(W, H, spp_gpu, seed=77, kernel=k))
    rs = ref_dof.scene("random_ned", seed=12345, wrap=True)
    r = rs.render(ref_dof.main_camera(W, H), W, H, spp_ref, seed=9, threads=8, want_moments=True)
    mean = r["sum"] / spp_ref
    var = np.maximum(r["sumsq"] / spp_ref - mean ** 2, 0.0)
    z2, flat = z2_statistic(img.astype(np.float64), spp_gpu, mean, var, spp_ref)
    print(f"dof/{kernel}: z2 {z2:.3f}, rays/path {st.rays / st.paths:.3f} vs {r['rays'] / r['paths']:.3f}")
    assert 0.8 < z2 < 1.2, f"z2 = {z2}"
    assert flat < 1e-3
    assert abs(st.rays / st.paths - r["rays"] / r["paths"]) < 0.02 * r["rays"] / r["paths"]
