#!/usr/bin/env python
"""Times rtow_render_accumulate_device for one scene / kernel / traversal (CUDA events on the launching stream).

  python tools/prof_render.py --scene cornell_smoke_ned --kernel 2 --width 1920 --height 1080 --spp 64 [--reps 2]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scene", default="random_ned")
    ap.add_argument("--kernel", type=int, default=0)
    ap.add_argument("--traverse", type=int, default=0)
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--reps", type=int, default=2)
    args = ap.parse_args()
    import torch
    import __graft_entry__ as entry
    pkg = entry.load_package()
    abi = pkg.abi
    host = pkg.HostScene(args.scene, seed=12345, wrap=True)
    scene = pkg.DeviceScene(host.desc, device=0)
    W, H = args.width, args.height
    cam = pkg.main_camera(W, H)
    p = abi.default_render_params(W, H, args.spp, seed=2024, kernel=args.kernel, traverse=args.traverse)
    acc = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda:0")
    stream = torch.cuda.current_stream()
    scene.accumulate_device(cam, p, acc.data_ptr(), stream.cuda_stream)
    torch.cuda.synchronize()
    a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(args.reps):
        scene.accumulate_device(cam, p, acc.data_ptr(), stream.cuda_stream)
    z.record(stream)
    torch.cuda.synchronize()
    st = scene.stream_stats(stream.cuda_stream)
    ms = a.elapsed_time(z) / args.reps
    print(f"{args.scene} kernel={args.kernel} {W}x{H}x{args.spp}: {ms:.2f} ms, {st.paths / ms / 1e3:.1f} Mpaths/s, "
          f"{st.rays / ms / 1e3:.1f} Mrays/s, rays/path {st.rays / max(st.paths, 1):.3f}, launches {st.launches}", flush=True)
    pr = scene.render_profile()
    if pr.sampled_iterations:
        n = pr.sampled_iterations
        print(f"  wavefront: {pr.iterations} iterations, pool {pr.pool_slots}, extend {pr.extend_ms_sum / n:.3f} ms/it, "
              f"shade {pr.shade_ms_sum / n:.3f} ms/it (sampled {n}); drain {pr.tail_iterations} iterations, {pr.tail_ms:.3f} ms (finish kernel: {pr.finish_launches}); {pr.polls} polls", flush=True)


if __name__ == "__main__":
    main()
