"""not-gpu: the N>1 path on CPU -- world_size 2 and 3 over gloo.  Each rank renders its sample range (here with the
oracle standing in for the kernel: the host-side sharding and reduce code is what is under test) and rank 0 must end
up with the full image's sums."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port_no, spp, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import __graft_entry__ as entry
    from oracle import bindings as ob
    pkg = entry.load_package()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    W, H = 40, 24
    host = pkg.HostScene("cornell_smoke_ned", seed=3)
    cam = pkg.main_camera(W, H)
    b, n = pkg.sharding.sample_range(spp, rank, world)
    acc = torch.zeros(H, W, 3, dtype=torch.float32)
    if n > 0:
        p = pkg.abi.default_render_params(W, H, spp, seed=17, sample_begin=b, sample_count=n)
        acc += torch.from_numpy(ob.PortOracle().render(host.desc, cam, p)["sum"])
    pkg.sharding.reduce_sums(acc, dst=0)
    if rank == 0:
        np.save(out_path, acc.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_sample_ranges_tile_exactly(pkg):
    for spp in (0, 1, 7, 16, 1024, 4096):
        for world in (1, 2, 3, 4, 8):
            seen = []
            for r in range(world):
                b, n = pkg.sharding.sample_range(spp, r, world)
                assert n >= 0 and abs(n - spp / world) < 1
                seen += list(range(b, b + n))
            assert seen == list(range(spp))
    with pytest.raises(ValueError):
        pkg.sharding.sample_range(8, 2, 2)


@pytest.mark.parametrize("world,spp", [(2, 8), (3, 5)])
def test_spp_sharded_render_sums_to_the_full_image(pkg, abi, port, tmp_path, world, spp):
    import torch.multiprocessing as mp
    out = str(tmp_path / "acc.npy")
    mp.spawn(_worker, args=(world, _free_port(), spp, out), nprocs=world, join=True)
    got = np.load(out)
    W, H = 40, 24
    host = pkg.HostScene("cornell_smoke_ned", seed=3)
    full = port.render(host.desc, pkg.main_camera(W, H), abi.default_render_params(W, H, spp, seed=17))["sum"]
    # same sample set, fp32 summation order differs (per-shard partial sums)
    assert np.allclose(got, full, rtol=2e-6, atol=1e-6)
    assert got.sum() > 0


def _band_worker(rank, world, port_no, H, W, band_rows, out_path):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import __graft_entry__ as entry
    pkg = entry.load_package()
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port_no))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = torch.arange(H * W * 3, dtype=torch.float32).reshape(H, W, 3)      # stands in for the rendered frame
    mine = pkg.sharding.band_ranges(H, rank, world, band_rows)
    bands = [full[y:y + n].clone() for y, n in mine]                          # what this rank would have rendered
    frame = torch.full((H, W, 3), -1.0)
    pkg.sharding.gather_bands(frame, mine, bands, W, dst=0, band_rows=band_rows)
    if rank == 0:
        np.save(out_path, frame.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_band_ranges_tile_exactly(pkg):
    for H in (1, 15, 16, 17, 270, 2160):
        for world in (1, 2, 3, 8):
            for rows in (1, 16, 64):
                seen = []
                for r in range(world):
                    for y, n in pkg.sharding.band_ranges(H, r, world, rows):
                        assert 1 <= n <= rows
                        seen += list(range(y, y + n))
                assert sorted(seen) == list(range(H))
    with pytest.raises(ValueError):
        pkg.sharding.band_ranges(10, 3, 3)


@pytest.mark.parametrize("world,H,rows", [(2, 37, 8), (3, 50, 16)])
def test_bands_gather_into_the_full_frame(pkg, tmp_path, world, H, rows):
    """Image-tile partitioning (SURVEY 8(e) partitioning 2): interleaved bands per rank, no reduce, one gather."""
    import torch.multiprocessing as mp
    out = str(tmp_path / "frame.npy")
    W = 11
    mp.spawn(_band_worker, args=(world, _free_port(), H, W, rows, out), nprocs=world, join=True)
    got = np.load(out)
    assert np.array_equal(got, np.arange(H * W * 3, dtype=np.float32).reshape(H, W, 3))
