"""Inputs of the golden vectors (tests/golden/*.npz), shared by the generator (tests/golden/make_golden.py,
which runs the reference) and by the tests (which run the oracle / the CUDA path).  Pure numpy; no
intersection or shading code."""
import numpy as np

# scenes whose world->hit() records are frozen; kw goes to the scene builders
HIT_SCENES = [
    ("random_ned", dict(wrap=True)),
    ("random", dict(wrap=False)),
    ("cornell_ned", dict(wrap=True)),
    ("cornell", dict(wrap=False)),
    ("cornell_smoke_ned", dict(wrap=True)),
    ("bvh_stress", dict(n=50000, seed=777)),
]

# converged-image moments of the reference's own pixel loop: (scene, kw, (W, H, spp))
IMAGE_CASES = [
    ("random_ned", dict(wrap=True), (64, 36, 2048)),
    ("cornell_ned", dict(wrap=True), (48, 27, 2048)),
    ("cornell_smoke_ned", dict(wrap=True), (48, 27, 2048)),
]


def render_threads_seed():
    return 8, 4242


def hit_rays(pkg, name, n=2048):
    """2048 primary rays of main()'s camera + 2048 incoherent rays (+ two pinhole grids for the BVH stress scene)."""
    rs = pkg.raysets
    cam = pkg.main_camera(400, 225)
    ext = 100.0 if name == "bvh_stress" else 600.0
    sets = [rs.camera_rays(cam, 400, 225, n, seed=11), rs.incoherent_rays(n, seed=12, extent=ext)]
    if name == "bvh_stress":
        sets.append(rs.coherent_rays(32))
        sets.append(rs.coherent_rays(32, origin=(0.0, 0.0, -120.0)))
    return np.concatenate(sets)


def probe_image():
    """A small RGB8 image for ImageTexture::value (texture.h:65-79)."""
    ny, nx = 37, 53
    y, x = np.mgrid[0:ny, 0:nx]
    return np.stack([(x * 5 + y) % 256, (y * 7 + 3) % 256, (x * y) % 256], axis=-1).astype(np.uint8)


def function_sites(n=512, seed=2026):
    rng = np.random.default_rng(seed)
    f32 = np.float32
    S = {}
    # AABB::hit: random boxes and rays, with the degenerate cases the slab test is sensitive to
    m = 2048
    lo = rng.uniform(-10, 10, (m, 3)).astype(f32)
    hi = lo + rng.uniform(0, 6, (m, 3)).astype(f32)
    rays = np.zeros(m, dtype=[("o", "<f4", 3), ("d", "<f4", 3), ("time", "<f4"), ("tmin", "<f4"), ("tmax", "<f4")])
    rays["o"] = rng.uniform(-15, 15, (m, 3)).astype(f32)
    rays["d"] = rng.uniform(-1, 1, (m, 3)).astype(f32)
    rays["tmin"] = 0.001
    rays["tmax"] = np.finfo(np.float32).max
    rays["d"][0::16, 0] = 0.0                       # zero component: +-inf slabs
    rays["d"][1::16, 1] = -0.0
    rays["o"][2::16, 2] = lo[2::16, 2]              # origin on a slab plane
    rays["d"][2::16, 2] = 0.0                       # ... with a zero component: 0/0 = NaN
    rays["tmax"][3::16] = 4.0
    rays["tmin"][4::16] = -100.0
    rays["o"][5::16] = (lo[5::16] + hi[5::16]) * f32(0.5)   # origin inside the box
    S["aabb_min"], S["aabb_max"], S["aabb_rays"] = lo, hi, rays
    # Perlin / checker / marble sites: near and far from the origin, negative coordinates included
    pts = np.concatenate([rng.uniform(-4, 4, (n // 2, 3)), rng.uniform(-1200, 1200, (n // 2, 3))]).astype(f32)
    S["points"] = pts
    u = rng.normal(size=(n, 3))
    S["unit_points"] = (u / np.linalg.norm(u, axis=1, keepdims=True)).astype(f32)
    uv = rng.uniform(0, 1, (n, 2)).astype(f32)
    uv[:8] = [(0, 0), (1, 1), (0, 1), (1, 0), (0.5, 0.5), (0.999999, 0.000001), (-0.25, 1.5), (1.5, -0.25)]
    S["uv"] = uv
    # reflect / refract / schlick
    v = rng.normal(size=(n, 3)).astype(f32) * rng.uniform(0.1, 900, (n, 1)).astype(f32)   # un-normalised, like the reference's rays
    nn = rng.normal(size=(n, 3))
    S["vec_v"] = v
    S["vec_n"] = (nn / np.linalg.norm(nn, axis=1, keepdims=True)).astype(f32)
    S["ni"] = np.where(rng.uniform(size=n) < 0.5, f32(1.5), f32(1.0 / 1.5)).astype(f32)
    S["cosine"] = rng.uniform(0, 1.5, n).astype(f32)
    S["ri"] = np.full(n, 1.5, dtype=f32)
    # tone curve input: linear radiance incl. 0, > 1 and a 20x emitter
    img = rng.uniform(0, 1.3, (16, 24, 3)).astype(f32)
    img[0, :4] = [(0, 0, 0), (1, 1, 1), (20, 0.5, 1e-6), (0.999, 1.0001, 3.0)]
    S["tone_image"] = img
    S["camera_sizes"] = [(400, 225), (1920, 1080), (800, 800)]
    S["film"] = [(0.0, 0.0), (1.0, 1.0), (0.5, 0.5), (0.25, 0.9), (0.731, 0.118)]
    return S
