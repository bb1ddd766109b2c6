"""Row N4 on the GPU, third slice: uhdrb200_jpeg_encode_{host,dev} — forward DCT, quantisation and Huffman coding on the
device, markers and byte stuffing on the host — against libjpeg-turbo's OWN output. The reference configures libjpeg with
jpeg_set_quality(q, TRUE), raw_data_in, JDCT_ISLOW, 2x2/1x1/1x1 sampling and the default tables
(lib/src/jpegencoderhelper.cpp:113-130) and pads partial MCUs with zero samples (:147-205). Pillow drives the same
libjpeg-turbo; it cannot take raw 4:2:0 planes, but a 4:4:4 YCbCr image whose chroma samples are replicated 2x2 is
downsampled by libjpeg back to exactly those planes ((4 v + 1 or 2) >> 2 == v), and an image padded with zeros to whole
MCUs differs from the reference's encode only in the SOF dimensions. So the bar is a BYTE-IDENTICAL file."""
import ctypes as C
import io

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth

pytestmark = pytest.mark.gpu
try:
    import PIL.Image as PIL
except ImportError:  # the committed golden CRCs still pin the encoder (test_encode_matches_committed_golden)
    PIL = None
needs_pillow = pytest.mark.skipif(PIL is None, reason="Pillow (libjpeg-turbo) not importable")


def planes_420(w, h, seed, noise=5.0):
    """Y (h x w), U, V (h/2 x w/2) with smooth structure, texture and a flat patch."""
    rng = np.random.Generator(np.random.PCG64(seed))
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    y = 128 + 90 * np.sin(xx / 19.0) * np.cos(yy / 23.0) + rng.normal(0, noise, (h, w))
    y[h // 4: h // 2, w // 4: w // 2] = 200
    cy, cx = np.mgrid[0:h // 2, 0:w // 2].astype(np.float32)
    u = 128 + 60 * np.sin(cx / 31.0 + 1.0) + rng.normal(0, noise / 2, (h // 2, w // 2))
    v = 128 + 60 * np.cos(cy / 27.0) + rng.normal(0, noise / 2, (h // 2, w // 2))
    return [np.clip(p, 0, 255).astype(np.uint8) for p in (y, u, v)]


def device_encode(lib, planes, quality, icc=None, strides=None, expect=0, cap=None):
    y = planes[0]
    h, w = y.shape
    grey = len(planes) == 1
    ls = (strides or (w, w // 2))[0]
    cs = (strides or (w, w // 2))[1]
    yb = np.full((h, ls), 0x33, np.uint8)
    yb[:, :w] = y
    img = abi.Image()
    img.data, img.width, img.height, img.luma_stride = yb.ctypes.data, w, h, ls
    keep = [yb]
    if grey:
        img.pixelFormat = abi.PIX_FMT_MONOCHROME
    else:
        cb = np.full((2 * (h // 2), cs), 0x44, np.uint8)
        cb[: h // 2, : w // 2] = planes[1]
        cb[h // 2:, : w // 2] = planes[2]
        img.chroma_data, img.chroma_stride, img.pixelFormat = cb.ctypes.data, cs, abi.PIX_FMT_YUV420
        keep.append(cb)
    cap = cap if cap is not None else w * h * 3 + 4096
    out = np.full(cap + 32, 0xA5, np.uint8)
    n = C.c_size_t(0)
    icc_buf = (C.c_ubyte * len(icc)).from_buffer_copy(icc) if icc else None
    st = lib.uhdrb200_jpeg_encode_host(C.byref(img), quality, icc_buf, len(icc) if icc else 0, out.ctypes.data, cap, C.byref(n))
    assert st == expect, (st, lib.uhdrb200_last_error())
    assert (out[cap:] == 0xA5).all()
    return bytes(out[: n.value]) if st == 0 else None


def pillow_encode(planes, quality, pad_to=None):
    """libjpeg-turbo's file for the same samples. pad_to = (W, H): samples beyond the image are zero, like compressYuv."""
    y = planes[0]
    h, w = y.shape
    W, H = pad_to or (w, h)
    if len(planes) == 1:
        full = np.zeros((H, W), np.uint8)
        full[:h, :w] = y
        im = PIL.fromarray(full, "L")
        kw = {}
    else:
        full = np.zeros((H, W, 3), np.uint8)
        full[:h, :w, 0] = y
        for k in (1, 2):
            c = np.zeros((H // 2, W // 2), np.uint8)
            c[: h // 2, : w // 2] = planes[k]
            full[..., k] = c.repeat(2, 0).repeat(2, 1)
        im = PIL.fromarray(full, "YCbCr")
        kw = {"subsampling": 2}
    b = io.BytesIO()
    im.save(b, "JPEG", quality=quality, **kw)
    return b.getvalue()


def with_dims(data, w, h):
    """The same file with other SOF0 dimensions."""
    d = bytearray(data)
    i = d.find(b"\xff\xc0")
    d[i + 5: i + 9] = bytes([h >> 8, h & 255, w >> 8, w & 255])
    return bytes(d)


@pytest.mark.parametrize("quality", [1, 10, 50, 75, 90, 95, 100])
@pytest.mark.parametrize("w,h", [(16, 16), (64, 48), (640, 368), (1920, 1088)])
@needs_pillow
def test_encode_420_byte_identical(uh, w, h, quality):
    planes = planes_420(w, h, w + quality)
    got, want = device_encode(uh.lib, planes, quality), pillow_encode(planes, quality)
    assert len(got) == len(want) and got == want, f"{w}x{h} q{quality}: {len(got)} vs {len(want)} bytes, first difference at " \
        f"{next((i for i, (a, b) in enumerate(zip(got, want)) if a != b), -1)}"


@pytest.mark.parametrize("quality", [30, 85, 100])
@pytest.mark.parametrize("w,h", [(8, 8), (320, 240), (960, 544)])
@needs_pillow
def test_encode_grey_byte_identical(uh, w, h, quality):
    """The gain map is one 8-bit plane (jpegr.cpp:1037-1050)."""
    plane = synth.natural_map(w, h, w + quality) if w >= 16 else np.arange(64, dtype=np.uint8).reshape(8, 8) * 3
    got, want = device_encode(uh.lib, [plane], quality), pillow_encode([plane], quality)
    assert got == want


@pytest.mark.parametrize("w,h", [(100, 60), (18, 10), (642, 362), (2, 2)])
@needs_pillow
def test_partial_mcus_are_zero_padded(uh, w, h):
    """Sizes that are not whole MCUs: compressYuv hands libjpeg zero rows / columns (jpegencoderhelper.cpp:147-205); the file
    equals libjpeg-turbo's for the zero-padded image except for the SOF dimensions."""
    planes = planes_420(w, h, w * 7 + h)
    W, H = -(-w // 16) * 16, -(-h // 16) * 16
    got = device_encode(uh.lib, planes, 88)
    assert got == with_dims(pillow_encode(planes, 88, pad_to=(W, H)), w, h)
    grey = [planes[0]]
    W8, H8 = -(-w // 8) * 8, -(-h // 8) * 8
    assert device_encode(uh.lib, grey, 70) == with_dims(pillow_encode(grey, 70, pad_to=(W8, H8)), w, h)


@needs_pillow
def test_strided_planes_and_icc(uh):
    w, h = 256, 144
    planes = planes_420(w, h, 3)
    plain = device_encode(uh.lib, planes, 92)
    assert device_encode(uh.lib, planes, 92, strides=(w + 18, w // 2 + 9)) == plain  # jpegr_test.cpp:1485-1582: strides do not matter
    icc = bytes(range(200)) * 3
    tagged = device_encode(uh.lib, planes, 92, icc=icc)
    app2 = b"\xff\xe2" + (len(icc) + 2).to_bytes(2, "big") + icc
    assert tagged == plain[:20] + app2 + plain[20:]  # jpeg_write_marker after SOI + JFIF APP0, jpegencoderhelper.cpp:100-102


@needs_pillow
def test_noise_quality_100_long_codes(uh):
    """i.i.d. noise at quality 100: every coefficient non-zero, the longest codes and sizes, runs of FF bytes to stuff."""
    rng = np.random.Generator(np.random.PCG64(11))
    planes = [rng.integers(0, 256, (368, 640), dtype=np.uint8), rng.integers(0, 256, (184, 320), dtype=np.uint8),
              rng.integers(0, 256, (184, 320), dtype=np.uint8)]
    got, want = device_encode(uh.lib, planes, 100), pillow_encode(planes, 100)
    assert got == want and b"\xff\x00" in got


def test_encode_matches_committed_golden(uh):
    """tests/golden/jpeg_encode_v1.json (made by make_jpeg_encode_golden.py with Pillow / libjpeg-turbo): size and CRC-32 of
    the files libjpeg-turbo writes for deterministic planes; needs no Pillow on the box."""
    import json
    import os
    import zlib

    from golden.make_jpeg_encode_golden import planes_of

    golden = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "jpeg_encode_v1.json")))
    for c in golden["cases"]:
        got = device_encode(uh.lib, planes_of(c["width"], c["height"], c["quality"], c["grey"]), c["quality"])
        assert (len(got), zlib.crc32(got)) == (c["bytes"], c["crc32"]), c


def test_error_codes(uh):
    lib = uh.lib
    planes = planes_420(64, 48, 1)
    n = C.c_size_t(0)
    out = np.zeros(1 << 16, np.uint8)
    assert lib.uhdrb200_jpeg_encode_host(None, 90, None, 0, out.ctypes.data, out.size, C.byref(n)) == abi.ERROR_BAD_PTR
    device_encode(lib, planes, 101, expect=abi.ERROR_INVALID_QUALITY_FACTOR)
    device_encode(lib, planes, 90, expect=abi.ERROR_BUFFER_TOO_SMALL, cap=300)   # not even the headers
    device_encode(lib, planes, 90, expect=abi.ERROR_BUFFER_TOO_SMALL, cap=900)   # headers fit, the scan does not
    img = abi.Image()
    img.data, img.chroma_data = planes[0].ctypes.data, planes[1].ctypes.data
    img.width, img.height, img.luma_stride, img.chroma_stride, img.pixelFormat = 64, 48, 60, 32, abi.PIX_FMT_YUV420
    assert lib.uhdrb200_jpeg_encode_host(C.byref(img), 90, None, 0, out.ctypes.data, out.size, C.byref(n)) == abi.ERROR_INVALID_STRIDE


@needs_pillow
def test_encode_dev_round_trip_on_a_side_stream(uh):
    """Planes in device memory (where API-0 leaves the SDR image) -> uhdrb200_jpeg_encode_dev on a non-default stream ->
    the same bytes as the host entry point; uhdrb200_jpeg_decode_dev of that file on the same stream gives the planes
    libjpeg-turbo decodes."""
    import torch as t

    from test_gpu_jpeg_decode import libjpeg_planes

    lib = uh.lib
    w, h = 1280, 720
    planes = planes_420(w, h, 21)
    want = device_encode(lib, planes, 90)
    d = t.from_numpy(np.concatenate([p.ravel() for p in planes])).cuda()
    img = abi.Image()
    img.data, img.chroma_data = d.data_ptr(), d.data_ptr() + w * h
    img.width, img.height, img.luma_stride, img.chroma_stride, img.pixelFormat = w, h, w, w // 2, abi.PIX_FMT_YUV420
    side = t.cuda.Stream()
    side.wait_stream(t.cuda.current_stream())
    out = np.zeros(w * h * 3, np.uint8)
    n = C.c_size_t(0)
    assert lib.uhdrb200_jpeg_encode_dev(C.byref(img), 90, None, 0, out.ctypes.data, out.size, C.byref(n), side.cuda_stream) == 0, lib.uhdrb200_last_error()
    got = bytes(out[: n.value])
    assert got == want
    d2 = t.zeros(w * h * 3 // 2, dtype=t.uint8, device="cuda")
    dimg = abi.Image()
    dimg.data = d2.data_ptr()
    buf = (C.c_ubyte * len(got)).from_buffer_copy(got)
    assert lib.uhdrb200_jpeg_decode_dev(buf, len(got), C.byref(dimg), side.cuda_stream) == 0, lib.uhdrb200_last_error()
    ref = np.concatenate([p.ravel() for p in libjpeg_planes(got)])
    assert np.array_equal(d2.cpu().numpy(), ref)
