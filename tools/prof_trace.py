#!/usr/bin/env python
"""Small driver for ncu: the closest-hit kernel (trace_kernel) on the config-5 scene, one launch per ray class.

  python tools/prof_trace.py [--spheres N] [--rays LOG2] [--traverse 2] [--reps R]

Prints Mrays/s per class (CUDA events on the launching stream).  Numbers printed under ncu are not bench values."""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spheres", type=int, default=1000000)
    ap.add_argument("--rays", type=int, default=20, help="log2 of the rays per class")
    ap.add_argument("--traverse", type=int, default=2)
    ap.add_argument("--reps", type=int, default=1)
    ap.add_argument("--scene", default="bvh_stress")
    args = ap.parse_args()
    import numpy as np
    import torch
    import __graft_entry__ as entry
    pkg = entry.load_package()
    abi = pkg.abi
    dev_t = torch.device("cuda", 0)
    stream = torch.cuda.current_stream()
    if args.scene == "bvh_stress":
        host = pkg.HostScene("bvh_stress", seed=12345, n=args.spheres)
        ext = 100.0
    else:
        host = pkg.HostScene(args.scene, seed=12345, wrap=True)
        ext = 600.0
    scene = pkg.DeviceScene(host.desc, device=0)
    n = 1 << args.rays
    side = int(round(n ** 0.5))
    sets = {"coherent": pkg.raysets.coherent_rays(side) if args.scene == "bvh_stress" else pkg.raysets.camera_rays(pkg.main_camera(1920, 1080), 1920, 1080, n, seed=3),
            "incoherent": pkg.raysets.incoherent_rays(n, seed=7, extent=ext)}
    if args.scene == "bvh_stress":
        sets["coherent_near"] = pkg.raysets.coherent_rays(side, origin=(0.0, 0.0, -120.0))
    for name, rays in sets.items():
        d_rays = torch.from_numpy(rays.view(np.uint8).reshape(len(rays), -1).copy()).to(dev_t)
        d_hits = torch.empty(len(rays), 48, dtype=torch.uint8, device=dev_t)
        scene.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr(), args.traverse, 0, stream.cuda_stream)
        torch.cuda.synchronize()
        a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(args.reps):
            scene.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr(), args.traverse, 0, stream.cuda_stream)
        z.record(stream)
        torch.cuda.synchronize()
        ms = a.elapsed_time(z) / args.reps
        nv, pt = scene.trace_counters(rays[:: max(1, len(rays) >> 16)], traverse=args.traverse)
        print(f"{name}: {len(rays)} rays, {ms:.3f} ms, {len(rays) / ms / 1e3:.1f} Mrays/s, node visits/ray {nv:.1f}, prim tests/ray {pt:.1f}", flush=True)


if __name__ == "__main__":
    main()
