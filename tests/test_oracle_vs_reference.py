"""not-gpu: the plain-C oracle and the host-side scene API against the REFERENCE ITSELF, live (oracle/_ref =
the reference's own headers compiled unmodified; built by oracle/Makefile where /root/reference exists, shipped
prebuilt to the GPU box).  Skipped when the prebuilt library is absent; tests/test_oracle_golden.py then still pins
the oracle through the committed vectors."""
import os

import numpy as np
import pytest

from conftest import SCENES, bits


def _earth(ob, ref, name):
    fn = "1_earth_8k.jpg" if name.endswith("_ned") else "earth1300x1300.jpg"
    return ref.decode_image(os.path.join(ob.REF_IMAGE_DIR, "input_images", fn))


def _same_arrays(a, b, label):
    for k in ("xforms", "prims", "objects", "nodes", "materials", "textures", "perlin"):
        assert a[k].dtype == b[k].dtype and a[k].tobytes() == b[k].tobytes(), f"{label}: {k} differ"
    assert a["header"] == b["header"], label
    assert len(a["images"]) == len(b["images"])
    for x, y in zip(a["images"], b["images"]):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("name,kw", SCENES + [("bvh_stress", dict(n=5000))], ids=[s[0] for s in SCENES] + ["bvh_stress"])
def test_host_scene_api_flattens_like_the_reference(pkg, ob, ref, name, kw):
    """include/rtow/scenes.h on include/rtow/scene_api.h vs scenes.h:29,122,220,355 on the reference's classes, both
    through the same flatten.h: every array byte-identical (primitive order, BVH topology and boxes from the
    reference's qsort build, wrapper chains, materials, textures, Perlin tables, decoded JPEG texels)."""
    rs = ref.scene(name, seed=12345, n=kw.get("n", 0), wrap=kw.get("wrap", True))
    # no image override: the library decodes input_images/*.jpg itself (include/rtow/jpeg_decode.h), the reference with
    # its stb_image (scenes.h:47,140), and the texel arrays must come out byte-identical
    hs = pkg.HostScene(name, seed=12345, n=kw.get("n", 0), wrap=kw.get("wrap", True))
    a, b = hs.arrays(), rs.arrays()
    if name.startswith("random"):
        assert len(a["images"]) == 1 and len(b["images"]) == 1, "the image texture must be present on both sides (no glass-sphere fallback)"
    _same_arrays(a, b, name)


def test_failed_image_load_gives_the_glass_sphere_like_the_reference(pkg, abi):
    """scenes.h:104-109,202-207: stbi_load returning NULL silently substitutes a Dielectric sphere."""
    h1, h2 = pkg.HostScene("random_ned", seed=12345, wrap=True), pkg.HostScene("random_ned", seed=12345, wrap=True, image=False)
    with_img, without = h1.arrays(), h2.arrays()          # views into the scenes' own storage: keep h1, h2 alive
    assert len(with_img["images"]) == 1 and len(without["images"]) == 0
    assert (with_img["textures"]["type"] == abi.TEX_IMAGE).sum() == 1 and (without["textures"]["type"] == abi.TEX_IMAGE).sum() == 0
    assert (without["materials"]["type"] == abi.MAT_DIELECTRIC).sum() == (with_img["materials"]["type"] == abi.MAT_DIELECTRIC).sum() + 1


@pytest.mark.parametrize("name,kw", SCENES + [("bvh_stress", dict(n=20000))], ids=[s[0] for s in SCENES] + ["bvh_stress"])
def test_oracle_closest_hit_is_bit_identical_to_world_hit(pkg, abi, ob, ref, port, name, kw):
    rs = ref.scene(name, seed=12345, n=kw.get("n", 0), wrap=kw.get("wrap", True))
    d = rs.desc
    cam = pkg.main_camera(400, 225)
    ext = 100.0 if name == "bvh_stress" else 600.0
    sets = [ob.camera_rays(cam, 400, 225, 1 << 14, seed=21), ob.incoherent_rays(1 << 14, seed=22, extent=ext),
            ob.incoherent_rays(1 << 12, seed=23, extent=ext / 20)]
    if name == "bvh_stress":
        sets += [ob.coherent_rays(96), ob.coherent_rays(96, origin=(0.0, 0.0, -120.0))]
    rays = np.concatenate(sets)
    rays["tmax"][5::7] = 900.0        # finite tmax and negative tmin are part of the hit() contract (hitable.h:21)
    rays["tmin"][3::11] = -5.0
    want, _ = rs.trace(rays, seed=3, threads=4)
    got = port.trace(d, rays, seed=3)
    mats = abi.desc_arrays(d)["materials"]
    medium = lambda h: (h["mat_id"] >= 0) & (mats["type"][np.maximum(h["mat_id"], 0)] == abi.MAT_ISOTROPIC)
    keep = ~(medium(want) | medium(got))      # free-flight draws come from different generators
    assert np.array_equal(got["mat_id"][keep], want["mat_id"][keep])
    known = keep & (want["prim_id"] >= 0)
    assert np.array_equal(got["prim_id"][known], want["prim_id"][known])
    hit = keep & (want["prim_id"] != -1)
    assert hit.mean() > 0.02
    for k in ("t", "p", "n"):
        assert np.array_equal(bits(got[k][hit]), bits(want[k][hit])), f"{name}: {k}"


def test_medium_hit_rate_matches_reference(pkg, abi, ob, ref, port):
    """ConstantMedium::hit (constantMedium.h:26-74) is stochastic: compare the fraction of rays that scatter inside
    the volumes and the mean free-flight distance (different generators, same distribution)."""
    rs = ref.scene("cornell_smoke_ned", seed=12345, wrap=True)
    d = rs.desc
    rays = ob.camera_rays(pkg.main_camera(400, 225), 400, 225, 1 << 16, seed=5)
    want, _ = rs.trace(rays, seed=9, threads=4)
    got = port.trace(d, rays, seed=9)
    mats = abi.desc_arrays(d)["materials"]
    medium = lambda h: (h["mat_id"] >= 0) & (mats["type"][np.maximum(h["mat_id"], 0)] == abi.MAT_ISOTROPIC)
    fw, fg = medium(want).mean(), medium(got).mean()
    assert fw > 0.02 and abs(fw - fg) < 4 * np.sqrt(fw / len(rays)) + 1e-3
    assert abs(want["t"][medium(want)].mean() - got["t"][medium(got)].mean()) < 0.05 * want["t"][medium(want)].mean()
    assert np.all(got["n"][medium(got)] == np.array([1, 0, 0], dtype=np.float32))    # constantMedium.h:66


def test_scatter_matches_reference(pkg, abi, ob, ref, port):
    """Material::scatter/emitted (material.h:34-148) on real hit records: attenuation and emission bit-identical; the
    scattered ray bit-identical where it is deterministic (Metal with fuzz 0; Dielectric = one of reflect/refract),
    inside the reference's unit ball around p+n for Lambertian."""
    rs = ref.scene("random_ned", seed=12345, wrap=True)
    d = rs.desc
    a = abi.desc_arrays(d)
    rays = np.concatenate([ob.camera_rays(pkg.main_camera(400, 225), 400, 225, 6000, seed=8),
                           ob.incoherent_rays(20000, seed=4, extent=400.0)])
    rays["o"][6000:, 2] -= 1000.0          # around the scene (its root Translate puts the ground ball at z = +1000... the spheres sit near z = 0)
    rays["o"][6000:] *= np.float32(0.5)
    hits = port.trace(d, rays, seed=1)
    seen = set()
    for i in np.flatnonzero(hits["prim_id"] >= 0):
        m = int(hits["mat_id"][i])
        mt = int(a["materials"]["type"][m])
        key = (mt, int(a["textures"]["type"][a["materials"]["tex_id"][m]]) if mt in (0, 3, 4) else -1)
        if key in seen and len(seen) >= 6:
            continue
        seen.add(key)
        ok_r, at_r, sc_r, em_r = rs.scatter(m, rays[i:i + 1], hits[i:i + 1], seed=5)
        ok_p, at_p, sc_p, em_p = port.scatter(d, hits[i:i + 1], rays[i:i + 1], seed=5)
        assert np.array_equal(bits(em_r), bits(em_p))
        if mt == abi.MAT_LAMBERTIAN:
            assert ok_r == ok_p == 1 and np.array_equal(bits(at_r), bits(at_p))
            for sc in (sc_r, sc_p):
                assert np.array_equal(bits(sc["o"]), bits(hits["p"][i]))
                assert np.linalg.norm(sc["d"] - hits["n"][i]) < 1.0 + 1e-5
            assert sc_r["time"] == sc_p["time"] == rays["time"][i]
        elif mt == abi.MAT_METAL:
            assert np.array_equal(bits(at_r), bits(at_p))
            if a["materials"]["f"][m][3] == 0:
                assert ok_r == ok_p and np.array_equal(bits(sc_r["d"]), bits(sc_p["d"]))
            assert sc_r["time"] == sc_p["time"] == 0
        elif mt == abi.MAT_DIELECTRIC:
            inside = float(np.dot(rays["d"][i], hits["n"][i])) > 0          # material.h:83-99: outward normal and n/n' by side
            nrm = -hits["n"][i:i + 1] if inside else hits["n"][i:i + 1]
            ni = a["materials"]["f"][m][0] if inside else a["materials"]["f"][m][1]
            refl, _, _ = ref.reflect_refract(rays["d"][i:i + 1], hits["n"][i:i + 1], np.array([ni], dtype=np.float32))
            _, refr, can = ref.reflect_refract(rays["d"][i:i + 1], nrm, np.array([ni], dtype=np.float32))
            assert ok_r == ok_p == 1 and np.array_equal(at_r, at_p)
            for sc in (sc_r, sc_p):      # Dielectric picks reflect or refract by one draw (material.h:109)
                assert sc["time"] == 0
                assert np.array_equal(bits(sc["d"]), bits(refl[0])) or (can[0] and np.array_equal(bits(sc["d"]), bits(refr[0])))
        elif mt == abi.MAT_DIFFUSE_LIGHT:
            assert ok_r == ok_p == 0 and em_r[0] == 20.0
    assert len(seen) >= 5, seen


def test_camera_frame_and_rays(pkg, ob, ref, port):
    for (W, H) in ((400, 225), (1920, 1080), (3840, 2160), (800, 800)):
        c = ref.main_camera(W, H)
        mine = pkg.main_camera(W, H)
        assert bytes(mine) == bytes(ref.camera_pod(c, private=True)) == bytes(ref.camera_pod(c, private=False))
        for s, t in ((0.0, 0.0), (0.3, 0.9), (1.0, 1.0)):
            a, b = ref.camera_ray(c, s, t, seed=2), port.camera_ray(mine, s, t, seed=2)
            assert np.array_equal(bits(a["o"]), bits(b["o"])) and np.array_equal(bits(a["d"]), bits(b["d"]))
            assert 0.0 <= b["time"] < 1.0


def test_per_function_probes_live(pkg, abi, ob, ref, port):
    rng = np.random.default_rng(99)
    pts = np.concatenate([rng.uniform(-3, 3, (2000, 3)), rng.uniform(-2000, 2000, (2000, 3))]).astype(np.float32)
    hs = pkg.HostScene("random_ned", seed=12345, wrap=True)
    assert np.array_equal(bits(port.perlin_turbulence(hs.desc, pts)), bits(ref.perlin_turbulence(pts)))
    assert np.array_equal(bits(port.perlin_noise(hs.desc, pts)), bits(ref.perlin_noise(pts)))
    u = rng.normal(size=(4000, 3))
    u = (u / np.linalg.norm(u, axis=1, keepdims=True)).astype(np.float32)
    assert np.array_equal(bits(port.sphere_uv(u)), bits(ref.sphere_uv(u)))
    img = rng.integers(0, 256, (23, 41, 3), dtype=np.uint8)
    uv = rng.uniform(-0.1, 1.1, (3000, 2)).astype(np.float32)
    hs2 = pkg.HostScene("random_ned", seed=12345, wrap=True, image=img)
    tid = int(np.flatnonzero(hs2.arrays()["textures"]["type"] == abi.TEX_IMAGE)[0])
    got = port.texture_value(hs2.desc, tid, uv, np.zeros((len(uv), 3), dtype=np.float32))
    assert np.array_equal(bits(got), bits(ref.image_texture_value(img, uv)))
    sums = rng.uniform(0, 40, (9, 13, 3)).astype(np.float32)
    mean, bgra = port.resolve(sums, 16)
    assert np.array_equal(bgra, ref.tonemap_bgra8(mean))


def test_defines_are_the_default_params(abi, ref):
    """defines.h:3-30 -> rtow_render_params defaults."""
    import ctypes as C
    vals = [C.c_int() for _ in range(6)] + [C.c_float(), C.c_float(), C.c_int()]
    ref.L.ref_defines(*[C.byref(v) for v in vals])
    p = abi.default_render_params(vals[0].value, vals[1].value, vals[2].value)
    assert (vals[0].value, vals[1].value, vals[2].value) == (800, 800, 64)
    assert p.max_depth == vals[4].value == 50 and p.global_illum_en == vals[5].value == 1
    assert p.sky_gain == vals[6].value and abs(p.ambient_gain - vals[7].value) < 1e-7 and vals[8].value == 0
