"""Host image IO (rtr_image_load / rtr_write_exr): SkyMap::new's EXR read (backgrounds.rs:58-80) and the fp32 output.

The reference reads EXR through the `exr` crate, which is not available here; the checker is OpenCV's OpenEXR
codec (an independent implementation): files written by cv2 in every supported compression / pixel type must load
to the same floats, and files written by rtr_write_exr must read back in cv2.  Small pre-made files under
tests/golden/images (written by tests/golden/make_images.py) are checked too, so the test does not depend on cv2
being able to write on the box.
"""
import os

import numpy as np
import pytest

os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
import raytracingrust_b200 as rt

GOLD = os.path.join(os.path.dirname(__file__), "golden", "images")


def _cv2():
    return pytest.importorskip("cv2")


def _picture(h, w, seed):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.stack([np.sin(xx * 0.11) * 0.5 + 0.5, (yy / max(h - 1, 1)) ** 2 * 40.0, np.cos(xx * 0.05 + yy * 0.07) * 3.0], -1)
    img = img.astype(np.float32) + rng.random((h, w, 3), dtype=np.float32) * 0.25
    img[h // 3, w // 2] = (5000.0, 0.0, 1e-4)  # a "sun" texel and a tiny value
    return img.astype(np.float32)


COMPRESSIONS = ["NO", "RLE", "ZIPS", "ZIP", "PIZ", "PXR24"]


@pytest.mark.parametrize("comp", COMPRESSIONS)
@pytest.mark.parametrize("ptype", ["HALF", "FLOAT"])
@pytest.mark.parametrize("shape", [(37, 53), (64, 128), (1, 1), (33, 7)])
def test_exr_read_matches_opencv(tmp_path, comp, ptype, shape):
    cv2 = _cv2()
    img = _picture(*shape, seed=shape[0] * 7 + shape[1])
    path = str(tmp_path / "t.exr")
    ok = cv2.imwrite(path, img[:, :, ::-1], [cv2.IMWRITE_EXR_TYPE, getattr(cv2, "IMWRITE_EXR_TYPE_" + ptype),
                                           cv2.IMWRITE_EXR_COMPRESSION, getattr(cv2, "IMWRITE_EXR_COMPRESSION_" + comp)])
    assert ok
    want = cv2.imread(path, cv2.IMREAD_UNCHANGED)[:, :, ::-1]
    got = rt.load_image(path)
    assert got.shape == want.shape and got.dtype == np.float32
    assert np.array_equal(got.view(np.uint32), np.ascontiguousarray(want, np.float32).view(np.uint32))
    if ptype == "FLOAT" and comp != "PXR24":  # lossless: the file holds the original floats
        assert np.array_equal(got.view(np.uint32), img.view(np.uint32))


def test_exr_with_alpha_and_grey(tmp_path):
    cv2 = _cv2()
    img = _picture(20, 30, 3)
    rgba = np.concatenate([img[:, :, ::-1], np.full((20, 30, 1), 0.5, np.float32)], -1)
    p = str(tmp_path / "a.exr")
    assert cv2.imwrite(p, rgba, [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_FLOAT])
    assert np.array_equal(rt.load_image(p), img)  # alpha dropped as SkyMap::new does (backgrounds.rs:66-69)
    g = str(tmp_path / "g.exr")
    assert cv2.imwrite(g, img[:, :, 0].copy(), [cv2.IMWRITE_EXR_TYPE, cv2.IMWRITE_EXR_TYPE_FLOAT])
    got = rt.load_image(g)
    assert np.array_equal(got[:, :, 0], img[:, :, 0]) and np.array_equal(got[:, :, 1], got[:, :, 2])


def test_hdr_read_matches_opencv(tmp_path):
    cv2 = _cv2()
    for shape in [(16, 40), (9, 5), (32, 300)]:  # width < 8: flat texels; otherwise run-length planes
        img = _picture(*shape, seed=5)
        p = str(tmp_path / "t.hdr")
        assert cv2.imwrite(p, img[:, :, ::-1])
        want = np.ascontiguousarray(cv2.imread(p, cv2.IMREAD_UNCHANGED)[:, :, ::-1], np.float32)
        got = rt.load_image(p)
        assert got.shape == want.shape
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("channels", [3, 4])
def test_exr_write_round_trip(tmp_path, channels):
    img = _picture(45, 70, 9)
    if channels == 4:
        img = np.concatenate([img, np.ones((45, 70, 1), np.float32)], -1)
    p = str(tmp_path / "w.exr")
    rt.write_exr(p, img)
    assert np.array_equal(rt.load_image(p).view(np.uint32), np.ascontiguousarray(img[:, :, :3]).view(np.uint32))
    cv2 = _cv2()
    back = cv2.imread(p, cv2.IMREAD_UNCHANGED)
    assert back.shape == (45, 70, channels)
    assert np.array_equal(back[:, :, 2::-1], img[:, :, :3])


def test_exr_write_incompressible_block_is_stored(tmp_path):
    rng = np.random.default_rng(1)
    img = rng.integers(0, 2**32, (16, 64, 3), dtype=np.uint32).view(np.float32)
    img = np.where(np.isfinite(img), img, 1.0).astype(np.float32)
    p = str(tmp_path / "n.exr")
    rt.write_exr(p, img)
    assert np.array_equal(rt.load_image(p).view(np.uint32), img.view(np.uint32))


def test_golden_image_files():
    want = np.load(os.path.join(GOLD, "expected.npy"))
    names = sorted(f for f in os.listdir(GOLD) if f.endswith((".exr", ".hdr")))
    assert len(names) >= 8
    for f in names:
        got = rt.load_image(os.path.join(GOLD, f))
        ref = np.load(os.path.join(GOLD, f + ".npy")) if os.path.exists(os.path.join(GOLD, f + ".npy")) else want
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), f


def test_skymap_from_path_uses_the_loader(tmp_path):
    img = _picture(8, 16, 2)
    p = str(tmp_path / "sky.exr")
    rt.write_exr(p, img)
    sky = rt.SkyMap(p)
    assert (sky.width, sky.height) == (16, 8)
    assert np.array_equal(sky.image.reshape(8, 16, 3), img)
    with pytest.raises(IOError):
        rt.SkyMap(str(tmp_path / "missing.exr"))


def test_unsupported_and_corrupt_files(tmp_path):
    p = tmp_path / "bad.exr"
    p.write_bytes(b"\x76\x2f\x31\x01\x02\x02\x00\x00")  # tiled flag
    with pytest.raises(rt.RtrError) as e:
        rt.load_image(str(p))
    assert "tiled" in str(e.value)
    p.write_bytes(b"not an image")
    with pytest.raises(rt.RtrError):
        rt.load_image(str(p))
    good = tmp_path / "good.exr"
    rt.write_exr(str(good), _picture(20, 20, 1))
    data = good.read_bytes()
    for cut in (10, 200, len(data) - 7):
        p.write_bytes(data[:cut])
        with pytest.raises(rt.RtrError):
            rt.load_image(str(p))


def test_exr_chunk_offset_near_2_64_is_rejected(tmp_path):
    """A crafted chunk offset close to 2^64 used to wrap the `offset + 8 > size` check and read out of bounds."""
    import struct
    good = tmp_path / "good.exr"
    rt.write_exr(str(good), _picture(20, 20, 1))  # ZIP, 16 lines per block: two chunks, a two-entry offset table
    data = bytearray(good.read_bytes())
    table = next(i for i in range(8, len(data) - 16) if struct.unpack_from("<Q", data, i)[0] == i + 16)
    for evil in (0xFFFFFFFFFFFFFFFC, 0xFFFFFFFFFFFFFFF8, len(data) - 4, 1 << 63):
        bad = bytearray(data)
        struct.pack_into("<Q", bad, table, evil)
        p = tmp_path / "evil.exr"
        p.write_bytes(bytes(bad))
        with pytest.raises(rt.RtrError):
            rt.load_image(str(p))
