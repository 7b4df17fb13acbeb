#!/usr/bin/env python
"""bench.py -- the reference's headline workload on N B200s of one node.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference ...                     the reference's CPU code, host cores

Workload (BASELINE.json configs[1]): the 'One Weekend' random-spheres scene main() renders
(Translate(randomScene_NED(),(0,0,1000)), camera of main.cpp:111-125), 1920x1080, 1024 spp,
depth 50.  One step = one full frame = W*H*spp camera paths.  N GPUs split the samples per pixel
(per-GPU Philox streams, same sample set for any N) and sum the float accumulators with one NCCL
reduce; total work is fixed, so scaling is "strong".

One JSON line on rank 0.  `value` = Mpaths/s with the scene resident in HBM (render kernel + NCCL
reduce + resolve, CUDA events on the launching stream, max over ranks).  `e2e` = the same metric
through the host-buffer entry point a caller of the reference's raytraceWorkerProcedure would use
(rtow_render: camera/params in, linear RGB + BGRA8 image out to pinned host memory).
`roofline` is the FP32-issue roofline BASELINE.json's north_star names (this path is neither HBM- nor
tensor-bound: SURVEY 8(d)); the HBM view is given beside it in `roofline_hbm`.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WIDTH, HEIGHT, SPP, DEPTH = 1920, 1080, 1024, 50
SCENE, SCENE_SEED, RENDER_SEED = "random_ned", 12345, 2024
WORKLOAD = "configs[1]: random-spheres scene (randomScene_NED + main.cpp camera), 1920x1080, 1024 spp, depth 50"

# Algorithmic work per unit, in fp32 lane-operations (DESIGN.md "Work model").
OPS_SLAB = 27          # ray-AABB slab test, division form (aabb.h:22-43)
OPS_NODE4 = 4 * 28 + 25  # fast traversal: 4 child boxes (6 plane fma + 5 exit bound + 3 slack + 12 inflated near/far + 2) + ordering network
OPS_PRIM = 24          # ray-sphere test, miss path (sphere.h:26-42); rect = 12
OPS_HIT = 42           # accepted hit: root, point, normal
OPS_SCATTER = 32       # Lambertian scatter incl. E[1.91] rejection rounds
OPS_NOISE = 1470       # Perlin marble: 7 octaves x ~210 (noise.h:38-49)
OPS_CHECKER = 10
OPS_IMAGE = 30
OPS_PATH = 60          # camera ray + RNG set-up + accumulation per path


def sm_count_and_clock(torch):
    p = torch.cuda.get_device_properties(0)
    mhz = None
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=clocks.max.sm", "--format=csv,noheader,nounits", "-i", "0"],
                             capture_output=True, text=True, timeout=20).stdout.strip().splitlines()
        mhz = float(out[0])
    except Exception:
        pass
    if mhz is None:
        try:
            mhz = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["sm_max_mhz"]
        except Exception:
            mhz = 1965.0
    return p.multi_processor_count, mhz


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._pump, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7 or not (t0 <= ts <= t1 + 0.3):
                continue
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "samples": len(sm), "reasons": sorted(reasons)}


def describe_images(images, scene):
    """What the image texture of the rendered scene is: both arms print it, so a mismatch cannot go unnoticed."""
    if not len(images):
        return "absent (glass-sphere fallback, scenes.h:202-207)" if scene.startswith("random") else "none in this scene"
    fn = "1_earth_8k.jpg" if scene.endswith("_ned") else "earth1300x1300.jpg"
    return ", ".join(f"{fn} {im.shape[1]}x{im.shape[0]}" for im in images)


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


# ---- the reference arm: the reference's own CPU code (oracle/_ref), else the C port ---------------
def reference_render_sample(threads, budget_s):
    """Times a bounded sample of the workload: the full 1920x1080 frame at `s` of the 1024 spp."""
    from oracle import bindings as ob
    if ob.have_ref():
        ref = ob.RefOracle()
        sc = ref.scene(SCENE, seed=SCENE_SEED, wrap=True)
        global REF_IMAGE_TEXTURE
        REF_IMAGE_TEXTURE = describe_images(sc.arrays()["images"], SCENE)
        cam = ref.main_camera(WIDTH, HEIGHT)
        probe = sc.render(cam, 480, 270, 1, seed=RENDER_SEED, threads=threads)      # calibration, untimed
        rate = probe["paths"] / max(probe["seconds"], 1e-6)
        spp = int(max(1, min(64, round(budget_s * rate / (WIDTH * HEIGHT)))))

        def step():
            r = sc.render(cam, WIDTH, HEIGHT, spp, seed=RENDER_SEED, threads=threads)
            return r["paths"], r["rays"], r["seconds"]
        return "reference", threads, f"full 1920x1080 frame at {spp} of 1024 spp per step, {threads} host threads (oracle/_ref: reference headers, g++ -O2 -ffp-contract=off)", step
    pkg = __import__("__graft_entry__").load_package()
    port = ob.PortOracle()
    host = pkg.HostScene(SCENE, seed=SCENE_SEED, wrap=True)
    cam = pkg.main_camera(WIDTH, HEIGHT)
    w, h = 480, 270

    def step():
        p = pkg.abi.default_render_params(w, h, 1, seed=RENDER_SEED, max_depth=DEPTH)
        t0 = time.perf_counter()
        r = port.render(host.desc, pkg.main_camera(w, h), p)
        return w * h, r["rays"], time.perf_counter() - t0
    return "port", 1, "480x270 at 1 spp per step, 1 thread (oracle/rtow_oracle.c; oracle/_ref not present)", step


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = host_threads()
    kind, cores, sample, step = reference_render_sample(threads, budget_s=8.0)
    for _ in range(args.warmup):
        step()
    paths = rays = 0
    secs = 0.0
    for _ in range(args.steps):
        p, r, s = step()
        paths += p; rays += r; secs += s
    value = paths / secs / 1e6
    line = {
        "impl": "reference", "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample, "image_texture": REF_IMAGE_TEXTURE},
        "mrays_per_s": rays / secs / 1e6,
        "cpu_baseline": {"value": value, "unit": "Mpaths/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mpaths/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), file=JSON_OUT, flush=True)


# ---- the product arm --------------------------------------------------------------------------
def run_product(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import __graft_entry__ as entry
    pkg = entry.load_package()
    abi = pkg.abi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or pkg.lib().rtow_device_count() < 1:
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # stdout carries the one JSON line only (NCCL's version banner goes there otherwise)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev_t = torch.device("cuda", local)
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream

    W, H, spp = args.width, args.height, args.spp
    host = pkg.HostScene(args.scene, seed=SCENE_SEED, wrap=True)      # image textures decoded by the library (include/rtow/jpeg_decode.h)
    image_texture = describe_images(host.arrays()["images"], args.scene)
    if args.scene.startswith("random") and image_texture.startswith("absent"):
        raise SystemExit("bench.py: input_images/1_earth_8k.jpg not found -- the scene would fall back to the glass sphere "
                         "(scenes.h:202-207) and no longer be the reference's; run __graft_entry__.build() or set RTOW_INPUT_IMAGES")
    scene = pkg.DeviceScene(host.desc, device=local)
    cam = pkg.main_camera(W, H)
    # spp sharding: rank g renders samples [b_g, b_g + n_g); the union is [0, spp) for any world size
    b = spp * rank // world
    n = spp * (rank + 1) // world - b
    my = abi.default_render_params(W, H, spp, seed=RENDER_SEED, max_depth=DEPTH, sample_begin=b, sample_count=n, traverse=args.traverse, kernel=args.kernel)
    full = abi.default_render_params(W, H, spp, seed=RENDER_SEED, max_depth=DEPTH, traverse=args.traverse, kernel=args.kernel)

    accum = torch.zeros(H, W, 3, dtype=torch.float32, device=dev_t)
    rgb = torch.empty(H, W, 3, dtype=torch.float32, device=dev_t)
    bgra = torch.empty(H, W, 4, dtype=torch.uint8, device=dev_t)
    h_rgb = torch.empty(H, W, 3, dtype=torch.float32).pin_memory()
    h_bgra = torch.empty(H, W, 4, dtype=torch.uint8).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev_t)   # > 126 MB L2

    cpu_pg = dist.new_group(backend="gloo") if world > 1 else None

    def barrier(on_cpu=False):
        # on_cpu: the waiting ranks block in a gloo barrier on the host, so that no NCCL kernel spins on their GPUs while rank 0
        # drives those GPUs itself (the N > 1 e2e leg)
        if world > 1:
            torch.cuda.synchronize()
            dist.barrier(group=cpu_pg) if on_cpu else dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        accum.zero_()
        scene.accumulate_device(cam, my, accum.data_ptr(), sptr)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            scene.resolve_device(accum.data_ptr(), W, H, spp, rgb.data_ptr(), bgra.data_ptr(), sptr)

    multi = None

    def step_e2e():
        # the call a user of the library makes: host buffers in and out.  N > 1: rtow_multi_render from ONE process (rank 0 of
        # this harness) drives all N GPUs -- one host thread per GPU, one reduce + resolve kernel over peer memory -- while the
        # other ranks wait at the barrier with their GPUs idle
        if world == 1:
            scene.render_into(cam, full, h_rgb.data_ptr(), h_bgra.data_ptr())   # rtow_render: synchronous, host buffers
        elif rank == 0:
            multi.render_into(cam, full, h_rgb.data_ptr(), h_bgra.data_ptr())   # rtow_multi_render: synchronous, host buffers

    def timed(step, steps, on_cpu=False):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        barrier(on_cpu)
        t0 = time.time()
        for a, z in ev:
            flush.fill_(1)            # L2 flush between timed iterations (not timed)
            barrier(on_cpu)
            a.record(stream)
            step()
            z.record(stream)
        barrier(on_cpu)
        t1 = time.time()
        ms = torch.tensor([sum(a.elapsed_time(z) for a, z in ev)], dtype=torch.float64, device=dev_t)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), t0, t1

    for _ in range(args.warmup):
        step_resident()
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    ms_total, t0, t1 = timed(step_resident, args.steps)
    clocks = sampler.stop(t0, t1) if sampler else None
    st = scene.stream_stats(sptr)                     # rays / kernel time of this rank's last launch
    kern_ms_last = st.kernel_ms
    rays_t = torch.tensor([float(st.rays)], dtype=torch.float64, device=dev_t)
    if world > 1:
        dist.all_reduce(rays_t, op=dist.ReduceOp.SUM)
    rays_per_step = float(rays_t.item())

    if world > 1 and rank == 0:
        multi = pkg.MultiDeviceScene(host.desc, list(range(world)))
    for _ in range(args.warmup):
        step_e2e()
    torch.cuda.synchronize()
    ms_e2e, _, _ = timed(step_e2e, args.steps, on_cpu=True)
    e2e_api = "rtow_render" if world == 1 else f"rtow_multi_render over {world} devices from one process ({multi.peer_mapped if multi else 0} peer-mapped)"
    if multi is not None:
        multi.close()

    # ---- the other render configs of BASELINE.json, same harness (all ranks take part) --------------------------
    extra = {}
    if not args.no_extra and (W, H, spp, args.scene) == (WIDTH, HEIGHT, SPP, SCENE):
        for key, (sc_name, w2, h2, spp2, what) in EXTRA_CONFIGS.items():
            extra[key] = render_block(pkg, abi, torch, dist, dev_t, stream, world, rank, local, sc_name, w2, h2, spp2, what, flush, args)

    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    paths_per_step = float(W) * H * spp
    ms_step = ms_total / args.steps
    value = paths_per_step / (ms_step * 1e-3) / 1e6
    e2e_value = paths_per_step / (ms_e2e / args.steps * 1e-3) / 1e6

    # ---- roofline of the dominant kernel (render_mega_kernel), measured live -------------------
    # work counters from a counting build of the same kernel on a 1-spp slice of the same frame
    cnt_params = abi.default_render_params(W, H, spp, seed=RENDER_SEED, max_depth=DEPTH, sample_begin=0, sample_count=min(4, spp), traverse=args.traverse)
    wc = scene.render_counters(cam, cnt_params)
    box_ops = OPS_NODE4 if wc.traverse == abi.TRAVERSE_FAST else OPS_SLAB
    ops_per_ray = (box_ops * wc.node_visits_per_ray + OPS_PRIM * wc.prim_tests_per_ray + OPS_HIT + OPS_SCATTER * wc.scatters_per_ray
                   + OPS_NOISE * wc.noise_evals_per_ray + OPS_CHECKER * wc.checker_evals_per_ray + OPS_IMAGE * wc.image_evals_per_ray)
    rays_per_path = wc.rays / max(1, wc.paths)
    my_paths = float(W) * H * n
    my_rays = float(st.rays)
    sms, mhz = sm_count_and_clock(torch)
    peak_tops = sms * 128 * mhz * 1e6 / 1e12
    prof = scene.render_profile()
    step_tops = (my_rays * ops_per_ray + my_paths * OPS_PATH) / (kern_ms_last * 1e-3) / 1e12
    if prof.kernel == abi.KERNEL_WAVEFRONT and prof.sampled_iterations > 0:
        # dominant kernel = wf_extend_kernel (world->hit for one bounce of the whole pool): traversal work only
        ext_ms = prof.extend_ms_sum / prof.sampled_iterations
        shade_ms = prof.shade_ms_sum / prof.sampled_iterations
        trav_ops_per_ray = box_ops * wc.node_visits_per_ray + OPS_PRIM * wc.prim_tests_per_ray
        achieved_tops = prof.pool_slots * trav_ops_per_ray / (ext_ms * 1e-3) / 1e12
        # list scenes and EXACT traversal run the reference walk in its own kernel (wavefront.cu run_wavefront)
        ext_name = "wf_extend_kernel" if wc.traverse == abi.TRAVERSE_FAST else ("wf_extend_list_kernel" if args.scene.startswith("cornell") else "wf_extend_exact_kernel")
        dom = {"drain": {"iterations_after_last_work_item": prof.tail_iterations, "ms": prof.tail_ms, "finish_kernel_launches": prof.finish_launches,
                         "host_polls_per_step": prof.polls},
               "kernel": ext_name, "kernel_ms": ext_ms, "launches_per_step": prof.iterations, "units_per_launch": prof.pool_slots,
               "ops_per_unit": trav_ops_per_ray, "share_of_step": ext_ms / (ext_ms + shade_ms), "shade_kernel_ms": shade_ms,
               "sampled_launches": prof.sampled_iterations}
        kernel_name = "wavefront"
    else:
        achieved_tops = step_tops
        dom = {"kernel": "render_mega_kernel", "kernel_ms": kern_ms_last, "launches_per_step": 1, "units_per_launch": my_rays,
               "ops_per_unit": ops_per_ray, "share_of_step": 1.0}
        kernel_name = "megakernel"
    # HBM view: algorithmic bytes = accumulator read+write once per launch (+ the scene, KBs)
    alg_bytes = W * H * 3 * 4 * 2
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        hbm_peak, hbm_src = float(peaks["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        hbm_peak, hbm_src = 6650.0, "fallback (B200_PROFILING.md)"
    hbm_ach = alg_bytes / (kern_ms_last * 1e-3) / 1e9
    hbm_note = "accumulator read+write only: this kernel is not HBM-bound"
    traffic = None
    if dom["kernel"].startswith("wf_extend"):
        # per launch and entry: ray record 32 B read, hit {t, prim} 8 B + queue entry 4 B written
        alg_bytes = prof.pool_slots * (32 + 8 + 4)
        hbm_ach = alg_bytes / (dom["kernel_ms"] * 1e-3) / 1e9
        hbm_note = dom["kernel"] + ": ray records read + hit/queue written per launch; the kernel is issue-bound, not HBM-bound"
        try:      # measured DRAM bytes per launch of the same kernel on the same scene and pool (one ncu --set full capture)
            t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[dom["kernel"]]
            if (W, H, args.scene) == (WIDTH, HEIGHT, SCENE) and prof.pool_slots == int(t.get("pool_slots", 1 << 20)):
                traffic = int(t["dram_bytes_per_launch"])
        except Exception:
            traffic = None

    # ---- config 5 beside it: BVH hit tests on 1 M random spheres (the north star's Gray/s figure) -
    bvh = None
    if not args.no_bvh:
        bvh = bvh_hit_bench(pkg, abi, torch, dev_t, stream, args)

    # ---- CPU baseline on this box's host cores, bounded sample ------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        kind, cores, sample, stepf = reference_render_sample(host_threads(), budget_s=10.0)
        p, r, s = stepf()
        cpu = {"value": p / s / 1e6, "unit": "Mpaths/s", "cores": cores, "kind": kind, "sample": sample, "mrays_per_s": r / s / 1e6}

    launches_per_step = int(st.launches) + 1   # this rank's render kernels + resolve_kernel (rank 0); NCCL's kernels are not ours
    line = {
        "metric": "Mpaths/s", "value": value, "unit": "Mpaths/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD if (W, H, spp, args.scene) == (WIDTH, HEIGHT, SPP, SCENE) else f"{args.scene} scene {W}x{H}, {spp} spp, depth {DEPTH}",
                   "image_texture": image_texture, "scene_seed": SCENE_SEED, "render_seed": RENDER_SEED, "parallelism": f"spp-sharded x{world}, NCCL sum-reduce of fp32 accumulators",
                   "kernel": kernel_name, "traverse": {1: "exact", 2: "fast"}[wc.traverse],
                   "l2": "flushed between timed iterations (256 MiB fill); scene is L2-resident by design"},
        "mrays_per_s": rays_per_step / (ms_step * 1e-3) / 1e6,
        "rays_per_path": rays_per_path,
        "e2e": {"value": e2e_value, "unit": "Mpaths/s", "h2d_bytes_per_step": C.sizeof(abi.Camera) + C.sizeof(abi.RenderParams),
                "d2h_bytes_per_step": W * H * 3 * 4 + W * H * 4, "ms_per_step": ms_e2e / args.steps, "api": e2e_api},
        "gpu_launches": launches_per_step * args.steps,
        "clocks": clocks,
        "roofline": dict({"bound": "fp32_issue", "achieved": achieved_tops, "peak": peak_tops, "unit": "Tlaneop/s", "frac": achieved_tops / peak_tops,
                          "traffic": traffic, "peak_source": f"{sms} SMs x 128 fp32 lanes x {mhz:.0f} MHz (clocks.max.sm)",
                          "whole_step_achieved": step_tops, "whole_step_frac": step_tops / peak_tops,
                          "ops_per_ray_incl_shading": ops_per_ray, "node_visits_per_ray": wc.node_visits_per_ray,
                          "prim_tests_per_ray": wc.prim_tests_per_ray, "noise_evals_per_ray": wc.noise_evals_per_ray,
                          "image_evals_per_ray": wc.image_evals_per_ray, "checker_evals_per_ray": wc.checker_evals_per_ray}, **dom),
        "roofline_hbm": {"bound": "hbm", "achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_ach / hbm_peak, "traffic": traffic,
                         "algorithmic_bytes_per_launch": alg_bytes, "peak_source": hbm_src, "note": hbm_note},
        "bvh_hit": bvh,
        "other_configs": extra,
        "cpu_baseline": cpu,
    }
    print(json.dumps(line), file=JSON_OUT, flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# BASELINE.json configs[2] and configs[3] beside the headline config: same harness, spp-sharded over the ranks.  The sample
# counts are bounded so that the default run stays within minutes (Mpaths/s does not depend on spp beyond the ~5 ms of
# per-frame fixed cost, which these counts already amortise); the full counts are 512 and 4096.
EXTRA_CONFIGS = {
    "configs[2]_textured": ("random_ned", 1920, 1080, 512, "configs[2]: textured scene (checker + Perlin noise + 2048x1024 JPEG earth image texture = randomScene_NED), 1920x1080, 512 spp"),
    "configs[3]_cornell_smoke": ("cornell_smoke_ned", 3840, 2160, 128, "configs[3]: Cornell box with constantMedium smoke volumes, 3840x2160, 128 of the 4096 spp per step (bounded sample)"),
}


def render_block(pkg, abi, torch, dist, dev_t, stream, world, rank, local, scene_name, W, H, spp, what, flush, args):
    host = pkg.HostScene(scene_name, seed=SCENE_SEED, wrap=True)
    scene = pkg.DeviceScene(host.desc, device=local)
    cam = pkg.main_camera(W, H)
    b = spp * rank // world
    n = spp * (rank + 1) // world - b
    my = abi.default_render_params(W, H, spp, seed=RENDER_SEED, max_depth=DEPTH, sample_begin=b, sample_count=n)
    accum = torch.zeros(H, W, 3, dtype=torch.float32, device=dev_t)
    rgb = torch.empty(H, W, 3, dtype=torch.float32, device=dev_t)
    bgra = torch.empty(H, W, 4, dtype=torch.uint8, device=dev_t)
    sptr = stream.cuda_stream

    def step():
        accum.zero_()
        scene.accumulate_device(cam, my, accum.data_ptr(), sptr)
        if world > 1:
            dist.reduce(accum, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            scene.resolve_device(accum.data_ptr(), W, H, spp, rgb.data_ptr(), bgra.data_ptr(), sptr)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    step()                                   # warm-up (allocations, first-touch)
    steps = 2
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for a, z in ev:
        flush.fill_(1)
        barrier()
        a.record(stream)
        step()
        z.record(stream)
    barrier()
    ms = torch.tensor([sum(a.elapsed_time(z) for a, z in ev)], dtype=torch.float64, device=dev_t)
    st = scene.stream_stats(sptr)
    rays = torch.tensor([float(st.rays)], dtype=torch.float64, device=dev_t)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    ms_step = float(ms.item()) / steps
    paths = float(W) * H * spp
    pr = scene.render_profile()
    out = {"workload": what, "image_texture": describe_images(host.arrays()["images"], scene_name), "value": paths / (ms_step * 1e-3) / 1e6, "unit": "Mpaths/s",
           "mrays_per_s": float(rays.item()) / (ms_step * 1e-3) / 1e6, "rays_per_path": float(rays.item()) / paths, "ms_per_step": ms_step, "steps": steps,
           "n_gpus": world, "mean_radiance": float(rgb.mean().item()) if rank == 0 else None}
    if pr.sampled_iterations:
        out["extend_kernel_ms"] = pr.extend_ms_sum / pr.sampled_iterations
        out["shade_kernel_ms"] = pr.shade_ms_sum / pr.sampled_iterations
        out["pool_slots"] = pr.pool_slots
    scene.close()
    return out


def bvh_hit_bench(pkg, abi, torch, dev_t, stream, args):
    """configs[4]: closest-hit-only harness on N random spheres, coherent + incoherent ray sets."""
    import numpy as np
    ob = pkg.raysets
    n_spheres, n_side, n_inc = args.bvh_spheres, 4096, 1 << 24      # 2^24 rays per class
    t0 = time.perf_counter()
    host = pkg.HostScene("bvh_stress", seed=SCENE_SEED, n=n_spheres)
    t1 = time.perf_counter()
    scene = pkg.DeviceScene(host.desc, device=dev_t.index)
    t2 = time.perf_counter()
    out = {"spheres": n_spheres, "host_build_s": t1 - t0, "upload_and_fast_build_s": t2 - t1}
    # the far pinhole (0,0,-300), the same grid from (0,0,-120) -- coherent rays outside the cancellation regime of
    # sphere.h:26-42, whose proven slack is what the far set pays for -- and the incoherent diffuse set
    sets = {"coherent": ob.coherent_rays(n_side), "coherent_near": ob.coherent_rays(n_side, origin=(0.0, 0.0, -120.0)),
            "incoherent": ob.incoherent_rays(n_inc, seed=7, extent=100.0)}
    for name, rays in sets.items():
        d_rays = torch.from_numpy(rays.view(np.uint8).reshape(len(rays), -1).copy()).to(dev_t)
        d_hits = torch.empty(len(rays), 48, dtype=torch.uint8, device=dev_t)
        for mode, tag in ((abi.TRAVERSE_FAST, "fast"), (abi.TRAVERSE_EXACT, "exact")):
            if tag == "exact" and args.bvh_skip_exact:
                continue
            reps = 3 if tag == "fast" else 1
            scene.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr(), mode, 0, stream.cuda_stream)   # warm-up
            torch.cuda.synchronize()
            a, z = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            for _ in range(reps):
                scene.trace_device(d_rays.data_ptr(), len(rays), d_hits.data_ptr(), mode, 0, stream.cuda_stream)
            z.record(stream)
            torch.cuda.synchronize()
            ms = a.elapsed_time(z) / reps
            out[f"{name}_{tag}_mrays_per_s"] = len(rays) / (ms * 1e-3) / 1e6
        # e2e: rtow_trace_rays with HOST buffers (pinned), the copies inside the timed region (CUDA events inside the call)
        h_rays = torch.from_numpy(rays.view(np.uint8).reshape(len(rays), -1).copy()).pin_memory()
        h_hits = torch.empty(len(rays), 48, dtype=torch.uint8).pin_memory()
        best = None
        for _ in range(3):
            st = abi.Stats()
            rc = pkg.lib().rtow_trace_rays(scene._h, h_rays.data_ptr(), len(rays), h_hits.data_ptr(), abi.TRAVERSE_FAST, 0, C.byref(st))
            if rc != abi.OK:
                raise RuntimeError(pkg.lib().rtow_last_error().decode())
            best = st.total_ms if best is None else min(best, st.total_ms)
        out[f"{name}_fast_e2e_mrays_per_s"] = len(rays) / (best * 1e-3) / 1e6
        out[f"{name}_e2e_h2d_bytes"], out[f"{name}_e2e_d2h_bytes"] = int(h_rays.numel()), int(h_hits.numel())
        assert bytes(h_hits[:4096].numpy().tobytes()) == bytes(d_hits[:4096].cpu().numpy().tobytes())
        del h_rays, h_hits
        hits = d_hits.cpu().numpy().view(abi.hit_dtype).reshape(-1)
        out[f"{name}_hit_fraction"] = float((hits["prim_id"] >= 0).mean())
        nv, pt = scene.trace_counters(rays[:: len(rays) >> 18], traverse=abi.TRAVERSE_FAST)    # every 64th ray: the whole frustum, not its first rows
        out[f"{name}_fast_node_visits_per_ray"], out[f"{name}_fast_prim_tests_per_ray"] = nv, pt
        # fractions of the FP32-issue roofline: counted work of this traversal, and the BVH-quality-independent
        # floor of 2*log2(N) slab tests + 1 sphere test = 1104 lane-ops per ray
        sms, mhz = sm_count_and_clock(torch)
        peak = sms * 128 * mhz * 1e6
        rate = out[f"{name}_fast_mrays_per_s"] * 1e6
        out[f"{name}_fast_counted_ops_per_ray"] = OPS_NODE4 * nv + OPS_PRIM * pt + OPS_HIT * out[f"{name}_hit_fraction"]
        out[f"{name}_fast_frac_fp32_issue_counted"] = rate * out[f"{name}_fast_counted_ops_per_ray"] / peak
        out[f"{name}_fast_frac_fp32_issue_floor1104"] = rate * 1104.0 / peak
    out["rays_per_class"] = {k: int(len(v)) for k, v in sets.items()}
    # the reference's own BvhNode::hit on the same spheres and ray classes, on this box's host cores (bounded sample); the GPU
    # hit records of the sampled rays are checked against it on the way (prim id exact, t to 1e-5 relative)
    if not args.no_cpu:
        from oracle import bindings as obind
        if obind.have_ref():
            threads = host_threads()
            ref = obind.RefOracle()
            t0 = time.perf_counter()
            rsc = ref.scene("bvh_stress", seed=SCENE_SEED, n=n_spheres)
            cpu = {"kind": "reference", "cores": threads, "reference_bvh_build_s": time.perf_counter() - t0,
                   "sample": f"{1 << 15} rays of each class (every {len(sets['coherent']) >> 15}th ray), world->hit() on the reference-built BvhNode, {threads} host threads"}
            for name, rays in sets.items():
                sub = np.ascontiguousarray(rays[:: len(rays) >> 15])
                ref_hits, secs = rsc.trace(sub, threads=threads)
                cpu[f"{name}_mrays_per_s"] = len(sub) / max(secs, 1e-9) / 1e6
                gpu_hits, _ = scene.trace(sub, traverse=abi.TRAVERSE_FAST)
                same = bool(np.array_equal(gpu_hits["prim_id"], ref_hits["prim_id"]))
                hit = ref_hits["prim_id"] >= 0
                terr = float(np.max(np.abs(gpu_hits["t"][hit] - ref_hits["t"][hit]) / np.abs(ref_hits["t"][hit]))) if hit.any() else 0.0
                cpu[f"{name}_gpu_matches_reference"] = same and terr <= 1e-5
            out["cpu_reference"] = cpu
    return out


JSON_OUT = sys.stdout
REF_IMAGE_TEXTURE = "decoded by the library (C port arm)"


def main():
    # stdout carries the one JSON line and nothing else: whatever a library writes to file descriptor 1 (NCCL's version banner
    # does, NCCL_DEBUG_FILE notwithstanding) is sent to stderr, and the line goes to a private duplicate of the real stdout
    global JSON_OUT
    sys.stdout.flush()
    JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--height", type=int, default=HEIGHT)
    ap.add_argument("--spp", type=int, default=SPP)
    ap.add_argument("--scene", default=SCENE, help="random_ned (configs[1], the default) | cornell_ned | cornell_smoke_ned (configs[3]'s scene)")
    ap.add_argument("--traverse", type=int, default=0, help="0 auto, 1 exact, 2 fast")
    ap.add_argument("--kernel", type=int, default=0, help="0 auto, 1 megakernel, 2 wavefront")
    ap.add_argument("--no-bvh", action="store_true", help="skip the config-5 BVH hit-test figures")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the configs[2] / configs[3] blocks")
    ap.add_argument("--bvh-spheres", type=int, default=1000000)
    ap.add_argument("--bvh-skip-exact", action="store_true")
    args = ap.parse_args()
    if args.scene != SCENE and not args.no_cpu:
        args.no_cpu = True          # the CPU legs time configs[1]; other scenes are scaling / tuning runs
    if args.impl == "reference":
        run_reference(args)
    else:
        run_product(args)


if __name__ == "__main__":
    main()
