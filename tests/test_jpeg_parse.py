"""Row N4, host side of the device JPEG decoder (no GPU): uhdrb200_jpeg_get_info = the marker parser of csrc/jpeg_huff.cu,
mirroring JpegDecoderHelper::getCompressedImageParameters and the layout checks of decompressImage
(lib/src/jpegdecoderhelper.cpp:254-258, 286-292, 339): sizes and formats of the reference's fixtures, and a clear refusal
(never a wrong decode) of everything the device decoder does not handle."""
import ctypes as C
import io

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, refdata
from libultrahdr_dev_b200.api import load_library

PIL = pytest.importorskip("PIL.Image")


def info(lib, data):
    w, h, f = C.c_size_t(0), C.c_size_t(0), C.c_int(-7)
    buf = (C.c_ubyte * max(1, len(data))).from_buffer_copy(bytes(data) or b"\0")
    st = lib.uhdrb200_jpeg_get_info(buf, len(data), C.byref(w), C.byref(h), C.byref(f))
    return st, w.value, h.value, f.value


def encode(arr, **kw):
    b = io.BytesIO()
    PIL.fromarray(arr).save(b, "JPEG", **kw)
    return b.getvalue()


@pytest.fixture(scope="module")
def lib():
    return load_library()


@pytest.mark.skipif(not all(refdata.jpeg_path(n) for n in refdata.JPEGS), reason="JPEG fixtures not staged")
def test_fixture_parameters(lib):
    want = {"jpeg_image.jpg": (1280, 720, abi.PIX_FMT_YUV420), "minnie-320x240-yuv.jpg": (320, 240, abi.PIX_FMT_YUV420),
            "minnie-320x240-y.jpg": (320, 240, abi.PIX_FMT_MONOCHROME)}
    for name, (w, h, f) in want.items():
        assert info(lib, open(refdata.jpeg_path(name), "rb").read()) == (0, w, h, f)


def test_generated_streams(lib):
    rng = np.random.Generator(np.random.PCG64(3))
    rgb = rng.integers(0, 256, (48, 64, 3), dtype=np.uint8)
    assert info(lib, encode(rgb, quality=90, subsampling=2)) == (0, 64, 48, abi.PIX_FMT_YUV420)
    assert info(lib, encode(rgb, quality=90, subsampling=2, optimize=True)) == (0, 64, 48, abi.PIX_FMT_YUV420)
    assert info(lib, encode(rgb[..., 0], quality=50)) == (0, 64, 48, abi.PIX_FMT_MONOCHROME)
    # what decompressImage refuses, or what the device decoder leaves to libjpeg
    st, *_ = info(lib, encode(rgb, quality=90, subsampling=0))  # 4:4:4: "decoding to YUV only supports 4:2:0 subsampling"
    assert st == abi.ERROR_UNSUPPORTED_FEATURE and b"4:2:0" in lib.uhdrb200_last_error()
    st, *_ = info(lib, encode(rgb, quality=90, subsampling=1))  # 4:2:2
    assert st == abi.ERROR_UNSUPPORTED_FEATURE
    st, *_ = info(lib, encode(rgb, quality=90, subsampling=2, progressive=True))
    assert st == abi.ERROR_UNSUPPORTED_FEATURE and b"progressive" in lib.uhdrb200_last_error()
    assert info(lib, encode(rgb, quality=90, subsampling=2, restart_marker_rows=1)) == (0, 64, 48, abi.PIX_FMT_YUV420)  # DRI is handled


def test_not_a_jpeg(lib):
    good = encode(np.zeros((16, 16, 3), np.uint8), quality=90, subsampling=2)
    assert info(lib, b"")[0] == abi.ERROR_DECODE
    assert info(lib, b"\x89PNG\r\n\x1a\n" + bytes(64))[0] == abi.ERROR_DECODE
    assert info(lib, good[:20])[0] == abi.ERROR_DECODE          # cut inside the headers
    assert info(lib, good[:2] + b"\xff\xe0\xff\xff" + good[2:])[0] == abi.ERROR_DECODE  # segment longer than the file
    assert lib.uhdrb200_jpeg_get_info(None, 10, None, None, None) == abi.ERROR_BAD_PTR
    buf = (C.c_ubyte * len(good)).from_buffer_copy(good)
    assert lib.uhdrb200_jpeg_get_info(buf, len(good), None, None, None) == 0  # outputs are optional


def test_mutated_headers_never_leave_the_buffer(lib):
    """What the reference's decoder fuzzer does to the whole decoder (fuzzer/ultrahdr_dec_fuzzer.cpp), here for the marker
    parser: byte flips, inserted bytes, cut files and over-long segment lengths in the header part of valid streams. Every
    call returns one of the documented codes; a stream that is accepted has a sane size and format. The mutated copy sits
    at the END of an exactly sized buffer followed by a guard page of different bytes, so a read past the end of the data
    shows up as a changed answer between two guard patterns."""
    rng = np.random.Generator(np.random.PCG64(99))
    rgb = rng.integers(0, 256, (40, 56, 3), dtype=np.uint8)
    seeds = [encode(rgb, quality=90, subsampling=2), encode(rgb, quality=60, subsampling=2, optimize=True),
             encode(rgb[..., 0], quality=80), encode(rgb, quality=90, subsampling=2, restart_marker_rows=1)]
    allowed = {0, abi.ERROR_DECODE, abi.ERROR_UNSUPPORTED_FEATURE, abi.ERROR_UNSUPPORTED_WIDTH_HEIGHT}
    accepted = 0
    for it in range(10000):
        d = bytearray(seeds[it % len(seeds)])
        head = min(len(d), 700)  # the markers live here
        kind = it % 5
        if kind == 0:
            for _ in range(int(rng.integers(1, 4))):
                d[int(rng.integers(2, head))] = int(rng.integers(0, 256))
        elif kind == 1:
            at = int(rng.integers(2, head))
            d[at:at] = bytes(rng.integers(0, 256, int(rng.integers(1, 9)), dtype=np.uint8))
        elif kind == 2:
            del d[int(rng.integers(2, len(d))):]
        elif kind == 3:  # a segment length that points far outside
            at = d.find(b"\xff\xdb")
            if at < 0:
                at = 2
            d[at + 2: at + 4] = bytes([int(rng.integers(0, 256)), int(rng.integers(0, 256))])
        else:
            at = int(rng.integers(2, head))
            d[at] = 0xFF
            d[at + 1] = int(rng.choice([0xC0, 0xC2, 0xC4, 0xDA, 0xDB, 0xDD, 0xD9, 0xE1]))
        answers = []
        for guard in (0x00, 0xFF):
            n = len(d)
            buf = (C.c_ubyte * (n + 4096)).from_buffer_copy(bytes(d) + bytes([guard]) * 4096)
            w, h, f = C.c_size_t(0), C.c_size_t(0), C.c_int(-7)
            st = lib.uhdrb200_jpeg_get_info(buf, n, C.byref(w), C.byref(h), C.byref(f))
            assert st in allowed, (it, st)
            answers.append((st, w.value, h.value, f.value))
        assert answers[0] == answers[1], (it, answers)  # nothing behind the last byte took part
        st, w, h, f = answers[0]
        if st == 0:
            accepted += 1
            assert 1 <= w <= 8192 and 1 <= h <= 8192 and f in (abi.PIX_FMT_YUV420, abi.PIX_FMT_MONOCHROME)
    assert 300 < accepted < 9700  # the mutations reach both sides


def test_size_limit(lib):
    big = encode(np.zeros((8, 8200), np.uint8), quality=50)  # wider than kMaxWidth = 8192 (jpegdecoderhelper.cpp:254-258)
    assert info(lib, big)[0] == abi.ERROR_UNSUPPORTED_WIDTH_HEIGHT


@pytest.mark.skipif(not refdata.jpeg_path("minnie-320x240-yuv.jpg"), reason="JPEG fixtures not staged")
def test_progressive_fixture_is_refused(lib):
    import os

    p = os.path.join(os.path.dirname(refdata.jpeg_path("minnie-320x240-yuv.jpg")), "minnie-320x240-yuv-icc.jpg")
    if not os.path.exists(p):
        p = "/root/reference/tests/data/minnie-320x240-yuv-icc.jpg"
    if not os.path.exists(p):
        pytest.skip("progressive fixture not available")
    assert info(lib, open(p, "rb").read())[0] == abi.ERROR_UNSUPPORTED_FEATURE  # SOF2


# ------------------------------------------------------------------------------ host side of the encoder (no GPU)
def _headers(lib, quality, grey, w, h, icc=b""):
    out = (C.c_ubyte * 70000)()
    icc_buf = (C.c_ubyte * len(icc)).from_buffer_copy(icc) if icc else None
    n = lib.uhdrb200_test_jpeg_headers(quality, 1 if grey else 0, w, h, icc_buf, len(icc), out, len(out))
    return bytes(out[:n])


@pytest.mark.parametrize("quality", [1, 25, 50, 51, 75, 90, 95, 100])
def test_encoder_headers_equal_libjpeg_turbo(lib, quality):
    """Markers, quantisation tables (jpeg_set_quality(q, TRUE): the Annex K tables scaled and clamped to 1..255) and Huffman
    tables (Annex K.3) in front of the scan: byte-identical to what libjpeg-turbo writes for the same settings."""
    rgb = np.zeros((32, 48, 3), np.uint8)
    for grey in (False, True):
        want = encode(rgb[..., 0] if grey else rgb, quality=quality, **({} if grey else {"subsampling": 2}))
        sos = want.find(b"\xff\xda")
        head = want[: sos + 2 + ((want[sos + 2] << 8) | want[sos + 3])]
        assert _headers(lib, quality, grey, 48, 32) == head


def test_encoder_icc_segment_and_limits(lib):
    plain = _headers(lib, 90, False, 64, 64)
    icc = bytes(range(256)) * 10
    tagged = _headers(lib, 90, False, 64, 64, icc)
    assert tagged == plain[:20] + b"\xff\xe2" + (len(icc) + 2).to_bytes(2, "big") + icc + plain[20:]
    out = (C.c_ubyte * 100)()
    assert lib.uhdrb200_test_jpeg_headers(90, 0, 64, 64, None, 0, out, 100) == 0  # does not fit
    assert _headers(lib, 90, False, 64, 64, bytes(65534)) == b""                  # one APP2 segment holds 65 533 bytes


def test_encoder_byte_stuffing(lib):
    rng = np.random.Generator(np.random.PCG64(4))
    for n in (0, 1, 15, 16, 17, 1000, 65537):
        src = rng.integers(0, 256, n, dtype=np.uint8)
        src[rng.random(n) < 0.1] = 0xFF  # dense FF bytes, runs of them included
        want = bytes(src).replace(b"\xff", b"\xff\x00")
        out = (C.c_ubyte * (2 * n + 32))()
        buf = (C.c_ubyte * max(1, n)).from_buffer_copy(bytes(src) or b"\0")
        got = lib.uhdrb200_test_jpeg_stuff(buf, n, out, len(out))
        assert bytes(out[:got]) == want
        if n > 17:
            assert lib.uhdrb200_test_jpeg_stuff(buf, n, out, len(want) - 1) == 0  # one byte short


def _destuff_model(src, align, max_intervals):
    """The rules of jpeg_destuff, byte by byte."""
    out, starts, marker, i, n = bytearray(), [0], 0, 0, len(src)
    while i < n:
        b = src[i]
        if b != 0xFF:
            out.append(b)
            i += 1
            continue
        m = src[i + 1] if i + 1 < n else 0xD9
        if m == 0x00:
            out.append(0xFF)
            i += 2
        elif m == 0xFF:
            i += 1
        elif 0xD0 <= m <= 0xD7 and len(starts) < max_intervals:
            out.extend(bytes(-len(out) % align))
            starts.append(len(out))
            i += 2
        else:
            marker = m
            break
    return bytes(out), starts, marker


def _destuff(lib, src, max_intervals):
    align = lib.uhdrb200_test_jpeg_subseq_bytes()
    n = len(src)
    out = (C.c_ubyte * (n + max_intervals * align + 64))()
    buf = (C.c_ubyte * max(1, n)).from_buffer_copy(bytes(src) or b"\0")
    marker, count = C.c_int(-1), C.c_uint(0)
    starts = (C.c_uint * max_intervals)()
    got = lib.uhdrb200_test_jpeg_destuff(buf, n, out, C.byref(marker), max_intervals, starts, C.byref(count))
    return bytes(out[:got]), list(starts[: count.value]), marker.value


def test_decoder_byte_stuffing_removal(lib):
    """The SSE2 loop of the decoder's host front end against the rules written out byte by byte: stuffed bytes, fill
    bytes, restart markers (padding to the subsequence size, interval offsets, more markers than intervals), the marker
    that ends the segment, FF as the very last byte, every length around the 16-byte vector size."""
    align = lib.uhdrb200_test_jpeg_subseq_bytes()
    assert align >= 32 and align % 4 == 0
    rng = np.random.Generator(np.random.PCG64(12))
    for n in list(range(0, 40)) + [127, 128, 129, 1000, 4097, 70001]:
        for dense in (0.0, 0.02, 0.3):
            src = rng.integers(0, 255, n, dtype=np.uint8)  # no FF yet
            at = np.nonzero(rng.random(n) < dense)[0]
            src = bytearray(bytes(src))
            for a in at:
                src[a] = 0xFF
                if a + 1 < n:  # mostly stuffing, some fill bytes and restart markers, never another marker
                    src[a + 1] = int(rng.choice([0x00, 0x00, 0x00, 0xFF, 0xD0 + int(rng.integers(0, 8))]))
            for max_intervals in (1, 3, 4096):
                want = _destuff_model(src, align, max_intervals)
                got = _destuff(lib, src, max_intervals)
                assert got == want, (n, dense, max_intervals)
    # a marker that is not RSTn ends the segment and is reported; so does an RSTn beyond the declared intervals
    body = bytes([1, 2, 3, 0xFF, 0x00, 4]) * 5
    for m in (0xD9, 0xC4, 0xDA, 0xE0):
        out, starts, marker = _destuff(lib, body + bytes([0xFF, m]) + body, 8)
        assert out == body.replace(b"\xff\x00", b"\xff") and marker == m and starts == [0]
    out, starts, marker = _destuff(lib, body + b"\xff\xd0" + body + b"\xff\xd1" + body, 2)
    assert marker == 0xD1 and starts == [0, align] and len(out) == align + len(body.replace(b"\xff\x00", b"\xff"))
