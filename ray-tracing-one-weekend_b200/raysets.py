"""Fixed ray sets of the closest-hit harness: pure numpy input generators,
no intersection code.  Used by bench.py (config 5) and, re-exported through oracle/bindings.py, by the tests."""
import numpy as np

from . import _abi as abi


# ---- ray sets of SURVEY section 8(d) C5 (fixed-ray harness) ---------------------------------
def coherent_rays(n_side, origin=(0.0, 0.0, -300.0), half_extent=0.45, time=0.5):
    """Pinhole at `origin` through an n_side x n_side grid on the plane z = origin.z + 1 (un-normalised d)."""
    g = (np.arange(n_side, dtype=np.float32) + np.float32(0.5)) / np.float32(n_side)
    xs = (g * 2 - 1) * np.float32(half_extent)
    X, Y = np.meshgrid(xs, xs, indexing="xy")
    rays = np.zeros(n_side * n_side, dtype=abi.ray_dtype)
    rays["o"] = np.asarray(origin, dtype=np.float32)
    rays["d"][:, 0] = X.ravel()
    rays["d"][:, 1] = Y.ravel()
    rays["d"][:, 2] = 1.0
    rays["time"] = time
    rays["tmin"] = 0.001
    rays["tmax"] = abi.FLT_MAX
    return rays


def incoherent_rays(n, seed=1, extent=100.0):
    """Origins ~U(-extent,extent)^3, directions uniform in the unit ball (rejection), random time."""
    rng = np.random.default_rng(seed)
    rays = np.zeros(n, dtype=abi.ray_dtype)
    rays["o"] = rng.uniform(-extent, extent, size=(n, 3)).astype(np.float32)
    d = np.zeros((n, 3), dtype=np.float32)
    todo = np.arange(n)
    while len(todo):
        c = rng.uniform(-1, 1, size=(len(todo), 3)).astype(np.float32)
        ok = (c * c).sum(1) < 1.0
        d[todo[ok]] = c[ok]
        todo = todo[~ok]
    rays["d"] = d
    rays["time"] = rng.uniform(0, 1, size=n).astype(np.float32)
    rays["tmin"] = 0.001
    rays["tmax"] = abi.FLT_MAX
    return rays


def camera_rays(cam, W, H, n, seed=3):
    """Primary rays of rtow_camera `cam` (DOF off) at n random film positions: camera.h:123-135 in numpy fp32."""
    rng = np.random.default_rng(seed)
    s = rng.uniform(0, 1, n).astype(np.float32)
    t = rng.uniform(0, 1, n).astype(np.float32)
    f = lambda a: np.asarray(a[:], dtype=np.float32)
    lr = np.float32(cam.lens_radius)
    off = f(cam.u) * lr + f(cam.v) * lr
    rays = np.zeros(n, dtype=abi.ray_dtype)
    rays["o"] = f(cam.origin) + off
    rays["d"] = f(cam.lower_left)[None, :] + s[:, None] * f(cam.horizontal)[None, :] + t[:, None] * f(cam.vertical)[None, :] \
        - f(cam.origin)[None, :] - off[None, :]
    rays["time"] = rng.uniform(0, 1, n).astype(np.float32)
    rays["tmin"] = 0.001
    rays["tmax"] = abi.FLT_MAX
    return rays
