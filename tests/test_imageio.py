"""Formats either side of the hot path: JPEG texture ingest, binary PPM and PNG output.

The JPEG decoder is checked byte for byte against libjpeg-turbo (through Pillow, test infrastructure only) on files
encoded here with every supported layout, and against the pinned decode of the reference's earth.jpg."""
import hashlib
import io
import os

import numpy as np
import pytest

import raytracingweekend_b200 as rtw
from conftest import GOLDEN

PIL = pytest.importorskip("PIL.Image")


def _test_image(w, h, seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    img = np.stack([(x * 255 // max(w - 1, 1)), (y * 255 // max(h - 1, 1)), ((x + y) * 7) % 256], axis=2).astype(np.float64)
    img += rng.normal(0, 20, img.shape)  # busy content: exercises long AC runs and every Huffman class
    return np.clip(img, 0, 255).astype(np.uint8)


@pytest.mark.parametrize("w,h", [(256, 256), (37, 21), (8, 8), (1, 1), (130, 67)])
@pytest.mark.parametrize("subsampling", [0, 1, 2])  # 4:4:4, 4:2:2, 4:2:0
@pytest.mark.parametrize("quality", [35, 90])
def test_baseline_jpeg_matches_libjpeg_byte_for_byte(w, h, subsampling, quality):
    src = PIL.fromarray(_test_image(w, h, w * 1000 + h))
    buf = io.BytesIO()
    src.save(buf, format="JPEG", quality=quality, subsampling=subsampling, progressive=False, optimize=(quality == 35))
    data = buf.getvalue()
    want = np.asarray(PIL.open(io.BytesIO(data)).convert("RGB"))
    got = rtw.decode_jpeg(data)
    assert got.shape == want.shape and np.array_equal(got, want)


def test_grayscale_restart_markers_and_errors():
    g = PIL.fromarray(_test_image(50, 40, 3)[..., 0])
    buf = io.BytesIO()
    g.save(buf, format="JPEG", quality=80)
    want = np.asarray(PIL.open(io.BytesIO(buf.getvalue())).convert("RGB"))
    assert np.array_equal(rtw.decode_jpeg(buf.getvalue()), want)
    # restart intervals (DRI + RSTn)
    c = PIL.fromarray(_test_image(96, 64, 4))
    buf = io.BytesIO()
    try:
        c.save(buf, format="JPEG", quality=75, subsampling=2, restart_marker_blocks=3)
    except TypeError:
        pytest.skip("this Pillow cannot write restart markers")
    want = np.asarray(PIL.open(io.BytesIO(buf.getvalue())).convert("RGB"))
    assert b"\xff\xdd" in buf.getvalue()
    assert np.array_equal(rtw.decode_jpeg(buf.getvalue()), want)
    # progressive files and garbage are refused, not mis-decoded
    buf = io.BytesIO()
    c.save(buf, format="JPEG", progressive=True)
    with pytest.raises(rtw._lib.RtwError):
        rtw.decode_jpeg(buf.getvalue())
    with pytest.raises(rtw._lib.RtwError):
        rtw.decode_jpeg(b"not a jpeg at all")
    with pytest.raises(rtw._lib.RtwError):
        rtw.decode_jpeg(buf.getvalue()[:100])


def test_corrupt_scan_headers_are_refused_and_fuzzed_files_never_crash():
    """A malformed SOS (Huffman table selectors beyond the four tables, zero-length segment) or DHT must fail with an error,
    never index outside the decoder's tables; random byte flips anywhere in the file either decode or fail cleanly."""
    c = PIL.fromarray(_test_image(48, 32, 5))
    buf = io.BytesIO()
    c.save(buf, format="JPEG", quality=70, subsampling=0)
    good = bytearray(buf.getvalue())
    sos = good.index(b"\xff\xda")
    ns = good[sos + 4]
    assert ns == 3
    for k in range(ns):  # each component's Td/Ta byte: selectors 4..15 do not exist
        bad = bytearray(good)
        bad[sos + 6 + 2 * k] = 0x7F
        with pytest.raises(rtw._lib.RtwError):
            rtw.decode_jpeg(bytes(bad))
    bad = bytearray(good)
    bad[sos + 2:sos + 4] = b"\x00\x02"  # SOS segment of length 2: no component count at all
    with pytest.raises(rtw._lib.RtwError):
        rtw.decode_jpeg(bytes(bad))
    dht = good.index(b"\xff\xc4")
    bad = bytearray(good)
    bad[dht + 4] = 0x25  # table class 2 does not exist
    with pytest.raises(rtw._lib.RtwError):
        rtw.decode_jpeg(bytes(bad))
    rng = np.random.default_rng(3)
    for _ in range(300):
        bad = bytearray(good)
        for pos in rng.integers(2, len(good), size=rng.integers(1, 4)):
            bad[pos] = rng.integers(0, 256)
        try:
            out = rtw.decode_jpeg(bytes(bad))
            assert out.ndim == 3 and out.shape[2] == 3
        except rtw._lib.RtwError:
            pass


def test_earth_jpg_decodes_to_the_pinned_texels():
    path = "/root/reference/RayTracingWeekend/earth.jpg"
    if not os.path.exists(path):
        pytest.skip("reference tree not present")
    tex = rtw.decode_jpeg(open(path, "rb").read())
    assert tex.shape == (256, 256, 3)
    assert hashlib.sha256(tex.tobytes()).hexdigest() == "da3cdb69263fdf2cf28811cab55d915e2ed89ad1a492d35f11e6087513f87e39"
    assert np.array_equal(tex.ravel(), np.fromfile(os.path.join(GOLDEN, "earth_256x256_rgb8.bin"), dtype=np.uint8))
    assert np.array_equal(tex, rtw.load_earth()) and np.array_equal(tex, rtw.load_earth(path))  # packaged texels == own decode


def test_png_and_binary_ppm_hold_the_same_pixels_as_the_ascii_ppm(tmp_path):
    rng = np.random.default_rng(5)
    nx, ny = 19, 11
    canvas = rng.random((ny, nx, 3))
    canvas[0, 0] = (0.0, 1.0, 0.5)
    p3, p6, png = str(tmp_path / "a.ppm"), str(tmp_path / "b.ppm"), str(tmp_path / "c.png")
    rtw.write_ppm(p3, canvas)
    rtw.write_ppm_binary(p6, canvas)
    rtw.write_png(png, canvas)
    tok = open(p3).read().split()
    assert tok[:4] == ["P3", str(nx), str(ny), "255"]
    ascii_px = np.array(tok[4:], dtype=np.int64).reshape(ny, nx, 3)
    want = (np.float32(255.99).astype(np.float64) * canvas[::-1]).astype(np.int64)  # :261 top row first, :268 int(255.99f*c)
    assert np.array_equal(ascii_px, want)
    raw = open(p6, "rb").read()
    head = f"P6\n{nx} {ny}\n255\n".encode()
    assert raw.startswith(head) and np.array_equal(np.frombuffer(raw[len(head):], dtype=np.uint8).reshape(ny, nx, 3), want)
    im = PIL.open(png)
    im.verify()  # CRCs
    assert np.array_equal(np.asarray(PIL.open(png).convert("RGB")), want)
    # multi-block zlib stream (> 65535 bytes of scanlines)
    big = rng.random((200, 200, 3))
    rtw.write_png(png, big)
    assert np.array_equal(np.asarray(PIL.open(png).convert("RGB")), (np.float32(255.99).astype(np.float64) * big[::-1]).astype(np.int64))
