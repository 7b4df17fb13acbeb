"""gpu: rtow_render_accumulate_device_noalloc (static __device__ scratch, no allocation at call time) against the
allocating entry rtow_render_accumulate_device, through the C ABI.  No counterpart in the reference
(include/rtow_b200.h).  Collected last on purpose
(file name).

The two entries run the same kernels on the same Philox streams and sum radiance in order-independent fixed point, so
the sums agree far inside the 2e-4 tolerance the render tests use; the check is repeated on a second call, because the
static array keeps the first call's contents (stale pool slots, queue entries, pixel sums)."""
import numpy as np
import pytest


def _sums(torch, fn, H, W):
    acc = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda:0")
    fn(acc.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return acc.cpu().numpy()


@pytest.mark.gpu
def test_noalloc_entry_matches_the_allocating_entry_twice_in_a_row(gpu, abi):
    import torch
    from conftest import earth_image
    calls = [
        # a BVH scene with every texture kind, more work items than the static pool holds (pool refills, drain tail)
        ("random_ned", 160, 90, 8, 3),
        # a list scene with media on the second call, smaller than the first: stale state of call 1 lies beyond what call 2 uses
        ("cornell_smoke_ned", 96, 54, 4, 11),
    ]
    for name, W, H, spp, seed in calls:
        host = gpu.HostScene(name, seed=12345, wrap=True, image=earth_image())
        dev = gpu.DeviceScene(host.desc, device=0)
        cam = gpu.main_camera(W, H)
        p = abi.default_render_params(W, H, spp, seed=seed)
        want = _sums(torch, lambda ptr, st: dev.accumulate_device(cam, p, ptr, st), H, W)
        for rep in range(2):
            got = _sums(torch, lambda ptr, st: dev.accumulate_device_noalloc(cam, p, ptr, st), H, W)
            err = float(np.abs(got - want).max())
            assert err <= 2e-4 * max(1.0, float(np.abs(want).max())), f"{name} call {rep}: max |diff| {err}"
        assert float(want.sum()) > 0.0
        dev.close()


@pytest.mark.gpu
def test_noalloc_entry_rejects_frames_beyond_its_static_scratch(gpu, abi):
    import torch
    host = gpu.HostScene("cornell_ned", seed=12345, wrap=True)
    dev = gpu.DeviceScene(host.desc, device=0)
    W, H = 1024, 512            # 524288 pixels > RTOW_NOALLOC_MAX_PIXELS (262144)
    cam = gpu.main_camera(W, H)
    acc = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda:0")
    with pytest.raises(gpu.RtowError) as e:
        dev.accumulate_device_noalloc(cam, abi.default_render_params(W, H, 1, seed=1), acc.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert e.value.code == abi.ERR_INVALID_ARG and "RTOW_NOALLOC_MAX_PIXELS" in str(e.value)
    # ... but a band of rows of the same frame fits (and equals the allocating entry's band)
    p = abi.default_render_params(W, H, 1, seed=1, row_begin=128, row_count=64)
    band = torch.zeros(64, W, 3, dtype=torch.float32, device="cuda:0")
    dev.accumulate_device_noalloc(cam, p, band.data_ptr(), torch.cuda.current_stream().cuda_stream)
    ref = torch.zeros_like(band)
    dev.accumulate_device(cam, p, ref.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert torch.allclose(band, ref, rtol=0, atol=2e-4 * max(1.0, float(ref.abs().max())))
    with pytest.raises(gpu.RtowError):
        dev.accumulate_device_noalloc(cam, abi.default_render_params(64, 36, 1, seed=1, kernel=abi.KERNEL_MEGA), band.data_ptr(), None)
