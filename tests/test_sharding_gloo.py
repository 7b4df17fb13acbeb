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
