#!/usr/bin/env python
"""Generates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref/libref_oracle.so = the reference's
own headers compiled unmodified from /root/reference, see oracle/Makefile).  Run in the build
container only (the GPU box has no /root/reference); the resulting files are committed.

    python tests/golden/make_golden.py

Every vector is an output of the reference's code for inputs that the tests can regenerate from seeds
(ray sets) or that are stored beside the outputs (per-function sites).  The reference has no golden
vectors or tests of its own, so these files are what pins the oracle (and through
it the CUDA path) where /root/reference is absent.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import __graft_entry__ as entry  # noqa: E402
from golden_spec import (HIT_SCENES, IMAGE_CASES, hit_rays, function_sites, probe_image, render_threads_seed)  # noqa: E402


def main():
    pkg = entry.load_package()
    abi = pkg.abi
    from oracle import bindings as ob
    ref = ob.RefOracle()

    # ---- closest-hit records: world->hit() of the reference on fixed ray sets --------------------
    for name, kw in HIT_SCENES:
        sc = ref.scene(name, seed=kw.get("seed", 12345), n=kw.get("n", 0), wrap=kw.get("wrap", True))
        rays = hit_rays(pkg, name)
        hits, _ = sc.trace(rays, seed=77)
        mats = abi.desc_arrays(sc.desc)["materials"]
        is_medium = (hits["mat_id"] >= 0) & (mats["type"][np.maximum(hits["mat_id"], 0)] == abi.MAT_ISOTROPIC)
        tag = name if "n" not in kw else f"{name}_{kw['n']}"
        np.savez_compressed(os.path.join(HERE, f"hits_{tag}.npz"), hits=hits, is_medium=is_medium,
                            rays_crc=np.uint32(np.frombuffer(rays.tobytes(), dtype=np.uint32).sum(dtype=np.uint64) & 0xffffffff))
        print(f"hits_{tag}: {len(rays)} rays, hit fraction {(hits['prim_id'] != -1).mean():.3f}, medium hits {int(is_medium.sum())}")

    # ---- per-function vectors --------------------------------------------------------------------
    S = function_sites()
    out = {}
    out["aabb"] = np.array([ref.aabb_hit(S["aabb_min"][i], S["aabb_max"][i], S["aabb_rays"][i:i + 1]) for i in range(len(S["aabb_rays"]))], dtype=np.uint8)
    out["perlin_noise"] = ref.perlin_noise(S["points"])
    out["perlin_turb"] = ref.perlin_turbulence(S["points"])
    out["sphere_uv"] = ref.sphere_uv(S["unit_points"])
    out["image_tex"] = ref.image_texture_value(probe_image(), S["uv"])
    refl, refr, ok = ref.reflect_refract(S["vec_v"], S["vec_n"], S["ni"])
    out["reflect"], out["refract"], out["refract_ok"] = refl, refr, ok
    out["schlick"] = ref.schlick(S["cosine"], S["ri"])
    out["tonemap"] = ref.tonemap_bgra8(S["tone_image"])
    # textures of the random scene: checker (children constant) and marble, by hit point
    sc = ref.scene("random_ned", seed=12345, wrap=True)
    texs = abi.desc_arrays(sc.desc)["textures"]
    for kind, key in ((abi.TEX_CHECKER, "tex_checker"), (abi.TEX_NOISE, "tex_noise")):
        tid = int(np.flatnonzero(texs["type"] == kind)[0])
        out[key] = sc.texture_value(tid, S["uv"][: len(S["points"])], S["points"])
    # Perlin tables as the reference's static initialisers built them (noise.h:58-116)
    out["perlin_tables"] = abi.desc_arrays(sc.desc)["perlin"].copy()
    # camera frame: private members of the reference Camera, and getRay() origins/directions
    cams, cam_rays = [], []
    for (W, H) in S["camera_sizes"]:
        c = ref.main_camera(W, H)
        pod = ref.camera_pod(c, private=True)
        cams.append(np.frombuffer(bytes(pod), dtype=np.uint8).copy())
        cam_rays.append(np.stack([ref.camera_ray(c, float(s), float(t), seed=1) for s, t in S["film"]]))
    out["camera_pods"] = np.stack(cams)
    out["camera_rays"] = np.stack(cam_rays)
    np.savez_compressed(os.path.join(HERE, "functions.npz"), **out)
    print("functions.npz:", {k: v.shape for k, v in out.items()})

    # ---- converged-image moments: the reference's pixel loop (main.cpp:932-956 around its own color()) ---
    threads, seed = render_threads_seed()
    for name, kw, (W, H, spp) in IMAGE_CASES:
        sc = ref.scene(name, seed=12345, wrap=True)
        cam = ref.main_camera(W, H)
        r = sc.render(cam, W, H, spp, seed=seed, threads=threads, want_moments=True)
        mean = (r["sum"] / spp).astype(np.float32)
        var = np.maximum(r["sumsq"] / spp - (r["sum"] / spp) ** 2, 0.0).astype(np.float32)   # per-sample variance
        np.savez_compressed(os.path.join(HERE, f"image_{name}.npz"), mean=mean, var=var, spp=np.int32(spp),
                            rays_per_path=np.float64(r["rays"] / r["paths"]))
        print(f"image_{name}: {W}x{H}x{spp}, {r['seconds']:.1f} s, rays/path {r['rays'] / r['paths']:.3f}, mean {mean.mean():.4f}")


if __name__ == "__main__":
    main()
