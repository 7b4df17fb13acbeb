"""not-gpu: the library's JPEG decoder (include/rtow/jpeg_decode.h, behind rtow::api::load_image) against the reference's
own stbi_load (oracle/_ref = the reference's vendored stb_image v2.22 compiled unmodified; scenes.h:47,140).  The decoded
texels are part of the flattened scene (ImageTexture::_data, texture.h:65-79), so they have to be byte-identical.

Three layers: (1) every JPEG the reference ships under input_images/ (where /root/reference exists), (2) sha256 of the two
files the builders actually load, pinned here so the check also runs where only the staged copies travel, (3) synthetic
files written with PIL in the flavours stb distinguishes: baseline / progressive, 4:4:4 / 4:2:2 / 4:2:0 chroma, grey,
restart intervals, sizes that are not a multiple of the MCU."""
import glob
import hashlib
import os

import numpy as np
import pytest

REF_DIR = "/root/reference/rayTracingOneWeekend/input_images"
# sha256 of the RGB8 texels stbi_load returns (made with oracle/_ref's ref_decode_image in this container)
PINNED = {
    "1_earth_8k.jpg": ((1024, 2048, 3), "45ede1416c89d1bffe600cc7835a77e052db626062e0d81ccf416d7bb533dc91"),
    "earth1300x1300.jpg": ((1300, 1300, 3), "7499d9a41305a8e298738a1435982d298ac4f1df861b1b11b0b437619e7e71fe"),
}


def _sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.skipif(not os.path.isdir(REF_DIR), reason="needs the reference checkout")
def test_every_reference_jpeg_decodes_like_stb_image(pkg, ref):
    files = sorted(glob.glob(os.path.join(REF_DIR, "*.jpg")))
    assert len(files) >= 2
    for fn in files:
        ours, theirs = pkg.decode_image(fn), ref.decode_image(fn)
        assert ours.shape == theirs.shape, fn
        assert np.array_equal(ours, theirs), f"{os.path.basename(fn)}: {(ours != theirs).sum()} bytes differ"


def test_builder_images_match_the_pinned_digests(pkg):
    d = pkg.image_dir()
    if d is None:
        pytest.skip("input_images/ not staged (build() copies them where the reference checkout exists)")
    for fn, (shape, digest) in PINNED.items():
        img = pkg.decode_image(os.path.join(d, fn))
        assert img.shape == shape
        assert _sha(img) == digest, fn


def _cases(tmp):
    from PIL import Image
    rng = np.random.default_rng(5)
    y, x = np.mgrid[0:97, 0:131]
    smooth = np.stack([127 + 120 * np.sin(x / 9.0), 127 + 120 * np.cos(y / 7.0), (x * y) % 256], axis=-1).astype(np.uint8)
    noisy = rng.integers(0, 256, size=(61, 83, 3), dtype=np.uint8)
    out = []
    for tag, img in (("smooth", smooth), ("noisy", noisy), ("tiny", smooth[:5, :3]), ("one", smooth[:1, :1])):
        for sub in (0, 1, 2):                      # 4:4:4, 4:2:2, 4:2:0
            for prog in (False, True):
                for q in (35, 92):
                    p = os.path.join(tmp, f"{tag}_{sub}_{int(prog)}_{q}.jpg")
                    Image.fromarray(img).save(p, quality=q, subsampling=sub, progressive=prog, optimize=prog)
                    out.append(p)
        p = os.path.join(tmp, f"{tag}_grey.jpg")
        Image.fromarray(img[..., 0]).save(p, quality=80)
        out.append(p)
    # restart intervals (cv2 writes DRI), baseline and progressive
    import cv2
    for prog in (0, 1):
        p = os.path.join(tmp, f"rst_{prog}.jpg")
        cv2.imwrite(p, smooth, [cv2.IMWRITE_JPEG_QUALITY, 70, cv2.IMWRITE_JPEG_RST_INTERVAL, 3, cv2.IMWRITE_JPEG_PROGRESSIVE, prog])
        out.append(p)
    return out


def test_synthetic_jpeg_flavours_decode_like_stb_image(pkg, ref, tmp_path):
    files = _cases(str(tmp_path))
    assert len(files) >= 54
    for fn in files:
        ours, theirs = pkg.decode_image(fn), ref.decode_image(fn)
        assert ours.shape == theirs.shape, fn
        assert np.array_equal(ours, theirs), f"{os.path.basename(fn)}: {(ours != theirs).sum()} bytes differ"


def test_decoder_rejects_what_is_not_a_jpeg(pkg, tmp_path):
    p = tmp_path / "junk.jpg"
    p.write_bytes(b"\xff\xd8\xff\xe0\x00\x10JFIF" + b"\x00" * 40)
    with pytest.raises(pkg.RtowError):
        pkg.decode_image(str(p))
    with pytest.raises(pkg.RtowError):
        pkg.decode_image(str(tmp_path / "missing.jpg"))
