#!/usr/bin/env python
"""torchrun --nproc-per-node N tools/check_multigpu.py : the N-GPU spp-sharded image (NCCL reduce of the fp32 sums)
equals the 1-GPU image of the same (seed, spp) up to fp32 summation order.  Prints one line on rank 0."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import numpy as np
    import torch
    import torch.distributed as dist
    import __graft_entry__ as entry
    pkg = entry.load_package()
    abi = pkg.abi
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H, spp = 480, 270, 64
    out = []
    for name in ("random_ned", "cornell_smoke_ned"):
        host = pkg.HostScene(name, seed=12345, wrap=True)
        scene = pkg.DeviceScene(host.desc, device=local)
        cam = pkg.main_camera(W, H)
        b, n = pkg.sharding.sample_range(spp, rank, world)
        acc = torch.zeros(H, W, 3, dtype=torch.float32, device=f"cuda:{local}")
        sptr = torch.cuda.current_stream().cuda_stream
        scene.accumulate_device(cam, abi.default_render_params(W, H, spp, seed=7, sample_begin=b, sample_count=n), acc.data_ptr(), sptr)
        pkg.sharding.reduce_sums(acc, dst=0)
        torch.cuda.synchronize()
        if rank == 0:
            full = torch.zeros_like(acc)
            scene.accumulate_device(cam, abi.default_render_params(W, H, spp, seed=7), full.data_ptr(), sptr)
            torch.cuda.synchronize()
            a, f = acc.cpu().numpy(), full.cpu().numpy()
            out.append(f"{name}: max|sharded-full| {np.abs(a - f).max():.3e} (sum scale {f.max():.1f}), allclose {np.allclose(a, f, rtol=2e-6, atol=1e-5)}")
    if rank == 0:
        print(f"world {world}: " + "; ".join(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
