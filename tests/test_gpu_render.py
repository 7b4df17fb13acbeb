"""-m gpu: the pixel/sample loop through the C ABI (rtow_render & friends) vs the oracle.
The GPU and the oracle draw the same Philox streams, so per-pixel sums agree to float rounding
(summation order, CUDA vs glibc sinf/atan2f last-bit differences): tolerance 2e-4 relative to the
image scale on >= 99.9 % of pixels, RMSE <= 1e-3 in linear radiance (BASELINE.json north_star)."""
import ctypes as C

import numpy as np
import pytest

from conftest import SCENES, earth_image

pytestmark = pytest.mark.gpu


def _render_pair(gpu, abi, port, name, kw, W, H, spp, seed=7, **pk):
    host = gpu.HostScene(name, seed=12345, image=earth_image(), **kw)
    dev = gpu.DeviceScene(host.desc)
    cam = gpu.main_camera(W, H)
    params = abi.default_render_params(W, H, spp, seed=seed, **pk)
    img, st = dev.render(cam, params)
    want = port.render(host.desc, cam, params)
    return img, st, want, (host, dev, cam, params)


@pytest.mark.parametrize("name,kw", SCENES)
@pytest.mark.parametrize("kernel", ["mega", "wavefront"])
def test_render_matches_oracle(gpu, abi, port, name, kw, kernel):
    W, H, spp = 96, 54, 8
    img, st, want, _ = _render_pair(gpu, abi, port, name, kw, W, H, spp,
                                    kernel=abi.KERNEL_MEGA if kernel == "mega" else abi.KERNEL_WAVEFRONT)
    ref = want["sum"] / np.float32(spp)
    assert st.paths == W * H * spp
    assert st.rays == want["rays"], "ray count differs: the GPU traced different paths than the oracle"
    scale = max(1.0, float(np.abs(ref).max()))
    close = np.abs(img - ref) <= 2e-4 * scale
    assert close.mean() >= 0.999, f"{name}: {100 * (1 - close.mean()):.3f}% of pixel channels off"
    rmse = float(np.sqrt(np.mean((img - ref) ** 2)))
    assert rmse <= 1e-3, f"{name}: RMSE {rmse}"


def test_camera_moves_without_reuploading_the_scene(gpu, abi, port):
    """the reference's fly-camera (main.cpp:485-642) only changes the camera; one DeviceScene, a fresh
    rtow_camera per frame (Camera::setLookFromPoint + setOrientationMatrix), every frame equal to the oracle's."""
    host = gpu.HostScene("cornell_ned", seed=12345)
    dev = gpu.DeviceScene(host.desc)
    W, H, spp = 80, 60, 4
    params = abi.default_render_params(W, H, spp, seed=11)
    frames = []
    for k, (eye, ang) in enumerate([((0, 0, 0), 0.0), ((40, 30, -20), 0.25), ((120, -60, 10), -0.4)]):
        c, s_ = np.cos(ang), np.sin(ang)
        m = np.array([[-c, s_, 0, 0], [s_, c, 0, 0], [0, 0, -1, 0], [0, 0, 0, 1]], dtype=np.float32)
        cam = gpu.oriented_camera(eye, m, 60.0, W / H)
        img, st = dev.render(cam, params)
        want = port.render(host.desc, cam, params)
        assert st.rays == want["rays"], f"frame {k}"
        ref = want["sum"] / np.float32(spp)
        assert np.abs(img - ref).max() <= 2e-4 * max(1.0, float(ref.max())), f"frame {k}"
        frames.append(img)
    assert not np.array_equal(frames[0], frames[1]) and not np.array_equal(frames[1], frames[2])


@pytest.mark.parametrize("kernel", [1, 2])
def test_depth_limit_and_no_global_illum(gpu, abi, port, kernel):
    for depth, gi in ((0, 1), (1, 1), (3, 0)):
        host = gpu.HostScene("cornell_smoke_ned", seed=1)
        dev = gpu.DeviceScene(host.desc)
        cam = gpu.main_camera(64, 36)
        params = abi.default_render_params(64, 36, 4, seed=5, max_depth=depth, kernel=kernel)
        params.global_illum_en = gi
        img, st = dev.render(cam, params)
        want = port.render(host.desc, cam, params)
        assert st.rays == want["rays"]
        assert np.abs(img - want["sum"] / np.float32(4)).max() <= 2e-4 * max(1.0, float(want["sum"].max()) / 4)


@pytest.mark.parametrize("W,H,spp", [(160, 90, 32), (37, 23, 70), (64, 36, 129)], ids=["160x90x32", "37x23x70", "64x36x129"])
def test_wavefront_equals_megakernel(gpu, abi, W, H, spp):
    """Same Philox streams, same path arithmetic: the two kernels differ only in the order of the per-pixel sum
    (fp32 shuffles vs 64-bit fixed point), and the wavefront image is deterministic although slots are handed out
    dynamically.  The sizes cover the work-item order of the wavefront path: a partial 64-sample block only, full blocks
    plus a partial one, a pixel count that is not a multiple of 32 (padding items).  (Slot regeneration -- more items than pool slots -- is
    exercised by test_converged_image_matches_reference_golden.)"""
    host = gpu.HostScene("random_ned", seed=12345, image=earth_image())
    dev = gpu.DeviceScene(host.desc)
    cam = gpu.main_camera(W, H)
    a, sa = dev.render(cam, abi.default_render_params(W, H, spp, seed=4, kernel=abi.KERNEL_MEGA))
    b, sb = dev.render(cam, abi.default_render_params(W, H, spp, seed=4, kernel=abi.KERNEL_WAVEFRONT))
    c, sc = dev.render(cam, abi.default_render_params(W, H, spp, seed=4, kernel=abi.KERNEL_WAVEFRONT))
    assert sa.rays == sb.rays == sc.rays and sa.paths == sb.paths == W * H * spp
    assert b.tobytes() == c.tobytes()
    assert np.allclose(a, b, rtol=2e-5, atol=2e-6)
    assert sb.launches > 3


def test_wavefront_fallback_to_the_reference_walk(gpu, abi, tmp_path):
    """Rays the FAST culling argument does not cover take the reference walk inside wf_extend_kernel.  A debug knob sends
    every 5th slot that way: the image must not change by a bit (FAST == EXACT hit for hit) and the ray counts must stay.
    Run in a child process because the knob is read once."""
    import os
    import subprocess
    import sys
    code = (
        "import sys, numpy as np\n"
        "sys.path.insert(0, %r)\n"
        "import __graft_entry__ as e\n"
        "pkg = e.load_package(); abi = pkg.abi\n"
        "sys.path.insert(0, %r)\n"
        "from conftest import earth_image\n"
        "host = pkg.HostScene('random_ned', seed=12345, image=earth_image())\n"
        "dev = pkg.DeviceScene(host.desc)\n"
        "img, st = dev.render(pkg.main_camera(96, 54), abi.default_render_params(96, 54, 40, seed=4, kernel=abi.KERNEL_WAVEFRONT))\n"
        "np.save(sys.argv[1], img); print(st.rays, st.paths)\n"
    ) % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for knob in ("0", "5"):
        path = str(tmp_path / f"img{knob}.npy")
        r = subprocess.run([sys.executable, "-c", code, path], capture_output=True, text=True, env=dict(os.environ, RTOW_WF_DEBUG_EXACT_EVERY=knob))
        assert r.returncode == 0, r.stdout + r.stderr
        outs.append((np.load(path), r.stdout.split()[-2:]))
    assert outs[0][1] == outs[1][1]                       # same ray and path counts
    assert outs[0][0].tobytes() == outs[1][0].tobytes()   # same image, bit for bit


@pytest.mark.parametrize("name", ["cornell_smoke_ned", "cornell_ned"])
def test_medium_boundary_evaluated_once_equals_the_two_pass_walk(gpu, abi, monkeypatch, name):
    """wf_extend_list_kernel evaluates a medium's boundary once for both queries of constantMedium.h:32,36 when it is a Box
    (a run of rectangles behind one chain) or one Sphere.  RTOW_LIST_BOUNDARY_ONCE=0 (read when the scene is uploaded) leaves
    the records unflagged, so every medium takes the generic two-pass walk: same image bit for bit, same ray counts."""
    host = gpu.HostScene(name, seed=12345, wrap=True)
    W, H = 128, 72
    cam = gpu.main_camera(W, H)
    p = abi.default_render_params(W, H, 24, seed=11, kernel=abi.KERNEL_WAVEFRONT)
    outs = []
    for knob in ("1", "0"):
        monkeypatch.setenv("RTOW_LIST_BOUNDARY_ONCE", knob)
        dev = gpu.DeviceScene(host.desc)
        img, st = dev.render(cam, p)
        outs.append((img.copy(), st.rays, st.paths))
        dev.close()
    assert outs[0][1:] == outs[1][1:]
    assert outs[0][0].tobytes() == outs[1][0].tobytes()
    assert outs[0][1] > 4 * W * H * 24            # long paths: the media and the boxes were walked


def test_render_is_deterministic_and_shard_invariant(gpu, abi):
    """spp sharding: the union of the shards' samples is the full image's sample set."""
    import torch
    host = gpu.HostScene("random_ned", seed=12345, image=earth_image())
    dev = gpu.DeviceScene(host.desc)
    W, H, spp = 128, 72, 16
    cam = gpu.main_camera(W, H)
    full = abi.default_render_params(W, H, spp, seed=99)
    a, _ = dev.render(cam, full)
    b, _ = dev.render(cam, full)
    assert a.tobytes() == b.tobytes()
    acc = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda:0")
    stream = torch.cuda.current_stream().cuda_stream
    for g in range(4):
        p = abi.default_render_params(W, H, spp, seed=99, sample_begin=g * 4, sample_count=4)
        dev.accumulate_device(cam, p, acc.data_ptr(), stream)
        st = dev.stream_stats(stream)
        assert st.paths == W * H * 4
    rgb = torch.empty_like(acc)
    bgra = torch.empty(H, W, 4, dtype=torch.uint8, device="cuda:0")
    dev.resolve_device(acc.data_ptr(), W, H, spp, rgb.data_ptr(), bgra.data_ptr(), stream)
    torch.cuda.synchronize()
    assert np.allclose(rgb.cpu().numpy(), a, rtol=1e-5, atol=1e-6)


def test_cpp_convenience_layer_renders(gpu, abi, tmp_path):
    """examples/render_api.cpp: scene built with the mirror scene API, flattened, uploaded once, two frames with a camera
    move in between, all through include/rtow/render.h.  The first frame equals the same render through ctypes."""
    import json
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.dirname(gpu.lib_path())
    exe = str(tmp_path / "render_api")
    r = subprocess.run(["g++", "-std=c++17", "-O1", os.path.join(root, "examples", "render_api.cpp"), "-I" + os.path.join(root, "include"),
                        "-L" + libdir, "-lrtow_b200", "-Wl,-rpath," + libdir, "-Wl,-rpath-link,/usr/local/cuda/lib64", "-o", exe],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    run = subprocess.run([exe, "64", "36", "4"], capture_output=True, text=True)
    assert run.returncode == 0, run.stderr
    st = json.loads(run.stdout.strip().splitlines()[-1])
    assert st["paths"] == [64 * 36 * 4] * 2 and min(st["rays"]) > 64 * 36 * 4 and st["mean"][0] != st["mean"][1]
    host = gpu.HostScene("cornell_ned", seed=12345)
    img, s0 = gpu.DeviceScene(host.desc).render(gpu.main_camera(64, 36), abi.default_render_params(64, 36, 4, seed=5))
    assert s0.rays == st["rays"][0] and abs(float(img.mean()) - st["mean"][0]) < 1e-5


def test_progressive_moments_checkpoint_and_resume(gpu, abi, port, tmp_path):
    """sums and sums of squares stay on the device across passes; two passes with a checkpoint in
    between are the same sample set as one pass; the second moment equals the oracle's (same Philox streams)."""
    import torch
    host = gpu.HostScene("cornell_smoke_ned", seed=12345)
    dev = gpu.DeviceScene(host.desc)
    W, H = 64, 36
    cam = gpu.main_camera(W, H)
    pr = gpu.ProgressiveRender(dev, cam, W, H, seed=21)
    st = pr.add(6)
    assert st.paths == W * H * 6
    pr.save(tmp_path / "ckpt.npz")
    del pr
    pr = gpu.ProgressiveRender.load(tmp_path / "ckpt.npz", dev, cam)
    assert pr.n == 6
    pr.add(10)
    params = abi.default_render_params(W, H, 16, seed=21)
    want = port.render(host.desc, cam, params, want_sumsq=True)
    got_sum, got_sq = pr.sum.cpu().numpy(), pr.sumsq.cpu().numpy()
    scale = max(1.0, float(want["sum"].max()))
    assert np.abs(got_sum - want["sum"]).max() <= 2e-4 * scale
    assert np.allclose(got_sq, want["sumsq"], rtol=2e-4, atol=1e-4 * max(1.0, float(want["sumsq"].max())))
    var = pr.variance_of_mean()
    assert var.shape == (H, W, 3) and np.all(var >= 0) and var.max() > 0
    assert np.allclose(pr.mean(), want["sum"] / np.float32(16), atol=2e-4 * scale)
    # the megakernel has no second-moment path: a loud error, not a silent zero image
    acc = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda:0")
    with pytest.raises(gpu.RtowError) as e:
        dev.accumulate_moments_device(cam, abi.default_render_params(W, H, 4, seed=1, kernel=abi.KERNEL_MEGA), acc.data_ptr(), acc.data_ptr(),
                                      torch.cuda.current_stream().cuda_stream)
    assert e.value.code == abi.ERR_UNSUPPORTED


@pytest.mark.parametrize("kernel", [1, 2], ids=["mega", "wavefront"])
def test_row_bands_tile_the_full_frame(gpu, abi, kernel):
    """SURVEY 8(e) partitioning 2: a band of rows rendered into a band-sized buffer holds exactly the rows of the full
    frame (pixels keep their full-frame Philox keys and camera positions); interleaved bands from sharding.band_ranges."""
    import torch
    host = gpu.HostScene("random_ned", seed=12345, image=earth_image())
    dev = gpu.DeviceScene(host.desc)
    W, H, spp = 72, 41, 6
    cam = gpu.main_camera(W, H)
    stream = torch.cuda.current_stream().cuda_stream
    full = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda:0")
    dev.accumulate_device(cam, abi.default_render_params(W, H, spp, seed=3, kernel=kernel), full.data_ptr(), stream)
    torch.cuda.synchronize()
    tiled = torch.zeros_like(full)
    paths = 0
    for rank in range(3):
        for y, n in gpu.sharding.band_ranges(H, rank, 3, band_rows=8):
            band = torch.zeros(n, W, 3, dtype=torch.float32, device="cuda:0")
            dev.accumulate_device(cam, abi.default_render_params(W, H, spp, seed=3, kernel=kernel, row_begin=y, row_count=n), band.data_ptr(), stream)
            paths += dev.stream_stats(stream).paths
            tiled[y:y + n] = band
    assert paths == W * H * spp
    assert torch.equal(tiled, full)
    # host-buffer entry point with a band: band-sized outputs
    img, st = dev.render(cam, abi.default_render_params(W, H, spp, seed=3, kernel=kernel, row_begin=16, row_count=9))
    assert img.shape == (9, W, 3) and st.paths == 9 * W * spp
    assert np.array_equal(img, (full[16:25] * np.float32(1.0 / spp)).cpu().numpy())
    with pytest.raises(gpu.RtowError):
        dev.render(cam, abi.default_render_params(W, H, spp, row_begin=40, row_count=5))


def test_bgra8_matches_reference_tone_curve(gpu, abi, ob):
    host = gpu.HostScene("cornell_ned", seed=3)
    dev = gpu.DeviceScene(host.desc)
    cam = gpu.main_camera(80, 45)
    rgb, bgra, _ = dev.render(cam, abi.default_render_params(80, 45, 4, seed=2), want_bgra=True)
    # main.cpp:955-989 in numpy: sqrt, 255.99 in double, truncate, clamp, B,G,R,0
    q = np.minimum((255.99 * np.sqrt(rgb).astype(np.float64)).astype(np.int64), 255).astype(np.uint8)
    want = np.stack([q[..., 2], q[..., 1], q[..., 0], np.zeros_like(q[..., 0])], axis=-1)
    assert np.array_equal(bgra, want)
    if ob.have_ref():
        assert np.array_equal(bgra, ob.RefOracle().tonemap_bgra8(rgb))


def test_bad_arguments_are_rejected(gpu, abi):
    host = gpu.HostScene("cornell_ned", seed=3)
    dev = gpu.DeviceScene(host.desc)
    cam = gpu.main_camera(8, 8)
    for bad in (dict(width=0), dict(spp=0), dict(sample_begin=5), dict(kernel=77), dict(traverse=9)):
        p = abi.default_render_params(8, 8, 4)
        for k, v in bad.items():
            setattr(p, k, v)
        with pytest.raises(gpu.RtowError):
            dev.render(cam, p)
    d = host.desc
    saved = d.abi_version
    d.abi_version = 99
    with pytest.raises(gpu.RtowError):
        gpu.DeviceScene(d)
    d.abi_version = saved


def test_headless_cli_writes_ppm_pfm_and_stats(gpu, abi, tmp_path):
    """The headless entry that replaces the Win32 shell: PPM carries the reference tone curve of the PFM's floats."""
    import json
    import os
    import subprocess
    cli = os.path.join(os.path.dirname(gpu.lib_path()), "rtow_cli")
    prefix = str(tmp_path / "f")
    out = subprocess.run([cli, "--scene", "cornell_smoke_ned", "--width", "64", "--height", "36", "--spp", "8", "--seed", "5", "--out", prefix,
                          "--exr"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    st = json.loads(out.stdout.strip().splitlines()[-1])
    assert st["paths"] == 64 * 36 * 8 and st["rays"] > st["paths"] and st["mean_radiance"] > 0
    pfm = open(prefix + ".pfm", "rb").read()
    assert pfm.startswith(b"PF\n64 36\n-1.0\n")
    rgb = np.frombuffer(pfm[len(b"PF\n64 36\n-1.0\n"):], dtype="<f4").reshape(36, 64, 3)[::-1]
    ppm = open(prefix + ".ppm", "rb").read()
    want = np.minimum((255.99 * np.sqrt(rgb).astype(np.float64)).astype(np.int64), 255).astype(np.uint8)
    assert ppm[len(b"P6\n64 36\n255\n"):] == want.tobytes()
    from test_abi_surface import _read_exr
    chans, exr = _read_exr(prefix + ".exr")
    assert chans == ["B", "G", "R"] and np.array_equal(exr, rgb)
    # the same frame through the Python binding
    host = gpu.HostScene("cornell_smoke_ned", seed=12345)
    img, _ = gpu.DeviceScene(host.desc).render(gpu.main_camera(64, 36), abi.default_render_params(64, 36, 8, seed=5))
    assert np.array_equal(img, rgb)
    # --gpus N: the same frame through rtow_multi_render (as many GPUs as the box has, at most 2 here)
    g = min(2, gpu.lib().rtow_device_count())
    out = subprocess.run([cli, "--scene", "cornell_smoke_ned", "--width", "64", "--height", "36", "--spp", "8", "--seed", "5", "--out", prefix + "m",
                          "--gpus", str(g)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    stm = json.loads(out.stdout.strip().splitlines()[-1])
    assert stm["gpus"] == g and stm["paths"] == st["paths"] and stm["rays"] == st["rays"]
    pfm = open(prefix + "m.pfm", "rb").read()
    rgbm = np.frombuffer(pfm[len(b"PF\n64 36\n-1.0\n"):], dtype="<f4").reshape(36, 64, 3)[::-1]
    assert np.abs(rgbm - rgb).max() <= 2e-5 * max(1.0, float(rgb.max()))
    # the random scene through the CLI carries the reference's image texture (found next to the library, whatever the CWD)
    if gpu.image_dir():
        out = subprocess.run([cli, "--scene", "random_ned", "--width", "32", "--height", "18", "--spp", "2", "--out", prefix + "r"],
                             capture_output=True, text=True, cwd=str(tmp_path))
        assert out.returncode == 0, out.stderr
        assert json.loads(out.stdout.strip().splitlines()[-1])["image_textures"] == 1


@pytest.mark.parametrize("kernel", ["mega", "wavefront"])
def test_converged_image_matches_reference_golden(gpu, abi, kernel):
    """BASELINE.json north_star: converged images must match the reference's CPU render.  tests/golden/image_*.npz holds the
    reference's own 2048-spp moments (mt19937_64 streams); the GPU renders 4096 spp (Philox).  Gate: variance-normalised
    z2 in (0.85, 1.15) -- i.e. the two are the same image up to Monte-Carlo noise -- and the RMSE that noise implies."""
    import os
    from golden_spec import IMAGE_CASES
    from test_oracle_golden import reference_earth, z2_statistic, load
    for name, kw, (W, H, ref_spp) in IMAGE_CASES:
        g = load(f"image_{name}.npz")
        host = gpu.HostScene(name, seed=12345, wrap=True, image=reference_earth())
        dev = gpu.DeviceScene(host.desc)
        spp = 4096
        p = abi.default_render_params(W, H, spp, seed=1234, kernel=abi.KERNEL_MEGA if kernel == "mega" else abi.KERNEL_WAVEFRONT)
        img, st = dev.render(gpu.main_camera(W, H), p)
        z2, flat = z2_statistic(img.astype(np.float64), spp, g["mean"].astype(np.float64), g["var"].astype(np.float64), int(g["spp"]))
        rmse = float(np.sqrt(np.mean((img - g["mean"]) ** 2)))
        expected = float(np.sqrt(np.mean(g["var"] * (1.0 / spp + 1.0 / int(g["spp"])))))
        print(f"{name}/{kernel}: z2 {z2:.3f}, RMSE {rmse:.4f} (noise floor {expected:.4f}), rays/path {st.rays / st.paths:.3f} vs {float(g['rays_per_path']):.3f}")
        assert 0.85 < z2 < 1.15, f"{name}: z2 = {z2}"
        assert flat < 1e-3 and rmse < 1.15 * expected
        assert abs(st.rays / st.paths - float(g["rays_per_path"])) < 0.01 * float(g["rays_per_path"])


def test_multi_gpu_render_through_the_c_abi(gpu, abi, port):
    """rtow_multi_render (one process, one host thread per shard, one reduce+resolve kernel over peer memory) equals the
    1-device rtow_render up to fp32 summation order, and the oracle.  With one GPU on the box both shards share it; with
    two or more the second shard sits on device 1 and is read over the peer mapping."""
    ndev = gpu.lib().rtow_device_count()
    host = gpu.HostScene("random_ned", seed=12345, wrap=True, image=earth_image())
    W, H, spp = 128, 72, 12
    cam = gpu.main_camera(W, H)
    p = abi.default_render_params(W, H, spp, seed=19)
    one, one_bgra, st1 = gpu.DeviceScene(host.desc).render(cam, p, want_bgra=True)
    for devices in ([0, 1 % ndev], [0, 0, 1 % ndev], None):
        m = gpu.MultiDeviceScene(host.desc, devices)
        assert m.shards == (len(devices) if devices else ndev)
        img, bgra, st = m.render(cam, p, want_bgra=True)
        assert st.paths == W * H * spp and st.rays == st1.rays
        assert np.abs(img - one).max() <= 2e-5 * max(1.0, float(one.max()))
        assert (np.abs(bgra.astype(int) - one_bgra.astype(int)) <= 1).all() and (bgra != one_bgra).mean() < 1e-3
        # a band of rows and a sample sub-range through the same entry
        pb = abi.default_render_params(W, H, spp, seed=19, row_begin=16, row_count=24, sample_begin=2, sample_count=7)
        band, _ = m.render(cam, pb)
        want, _ = gpu.DeviceScene(host.desc).render(cam, pb)
        assert band.shape == (24, W, 3) and np.abs(band - want).max() <= 2e-5 * max(1.0, float(want.max()))
        m.close()
    ref = port.render(host.desc, cam, p)["sum"] / np.float32(spp)
    assert np.abs(img - ref).max() <= 2e-4 * max(1.0, float(ref.max()))
    with pytest.raises(gpu.RtowError):
        gpu.MultiDeviceScene(host.desc, [0, ndev + 3])
