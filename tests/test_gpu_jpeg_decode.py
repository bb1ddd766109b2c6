"""Row N4 on the GPU, second slice: uhdrb200_jpeg_decode_{host,dev} — marker parsing and byte-stuffing removal on the host,
parallel Huffman decoding (csrc/jpeg_huff.cu) and the ISLOW IDCT (csrc/jpeg_idct.cu) on the device — against
libjpeg-turbo's OWN output: the committed CRCs of the reference's JPEG fixtures (tests/golden/jpeg_planes_v1.json) and,
through Pillow, libjpeg-turbo directly on the fixtures and on generated images of many sizes, qualities and table sets.
The bar is byte-identical planes (entropy decoding and the integer IDCT are exact)."""
import ctypes as C
import io
import json
import os
import zlib

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, refdata, synth

import parity
from test_jpeg_idct import need_fixtures

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FLT_MAX = 3.4028234663852886e38


def decode_host(lib, data, expect=0):
    w, h, f = C.c_size_t(0), C.c_size_t(0), C.c_int(-1)
    buf = (C.c_ubyte * len(data)).from_buffer_copy(data)
    assert lib.uhdrb200_jpeg_get_info(buf, len(data), C.byref(w), C.byref(h), C.byref(f)) == 0, lib.uhdrb200_last_error()
    w, h = w.value, h.value
    n = w * h if f.value == abi.PIX_FMT_MONOCHROME else w * h + 2 * (w // 2) * (h // 2)
    out = np.full(n + 64, 0xA5, np.uint8)  # guard band behind the planes
    img = abi.Image()
    img.data = out.ctypes.data
    st = lib.uhdrb200_jpeg_decode_host(buf, len(data), C.byref(img))
    assert st == expect, lib.uhdrb200_last_error()
    assert (out[n:] == 0xA5).all()
    if st:
        return None
    assert (img.width, img.height, img.luma_stride, img.pixelFormat) == (w, h, w, f.value)
    planes = [out[: w * h].reshape(h, w)]
    if f.value == abi.PIX_FMT_YUV420:
        cw, ch = w // 2, h // 2
        assert img.chroma_data == img.data + w * h and img.chroma_stride == cw
        planes += [out[w * h: w * h + cw * ch].reshape(ch, cw), out[w * h + cw * ch: n].reshape(ch, cw)]
    return planes


def libjpeg_planes(data):
    from golden.make_jpeg_golden import libjpeg_planes as f

    return f(io.BytesIO(data))


def encode(arr, **kw):
    from PIL import Image

    b = io.BytesIO()
    Image.fromarray(arr).save(b, "JPEG", **kw)
    return b.getvalue()


def scene(w, h, seed, noise=6.0):
    """A smooth colour field with texture and hard edges, as RGB."""
    rng = np.random.Generator(np.random.PCG64(seed))
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.stack([127 + 100 * np.sin(xx / (17.0 + 5 * c) + c) * np.cos(yy / (23.0 - 3 * c)) for c in range(3)], axis=-1)
    img += rng.normal(0, noise, img.shape)
    img[h // 3: h // 3 + max(1, h // 7), w // 4: w // 2] = rng.integers(0, 256, 3)  # a flat rectangle: long runs of EOB
    return np.clip(img, 0, 255).astype(np.uint8)


# ------------------------------------------------------------------------------ the reference's fixtures
@need_fixtures
@pytest.mark.parametrize("name", refdata.JPEGS)
def test_fixture_decode_equals_libjpeg_turbo(uh, name):
    data = open(refdata.jpeg_path(name), "rb").read()
    planes = decode_host(uh.lib, data)
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "jpeg_planes_v1.json")))["files"][name]
    assert len(planes) == len(golden)
    for p, g in zip(planes, golden):
        assert list(p.shape) == g["shape"] and zlib.crc32(p.tobytes()) == g["crc32"]
    try:
        ref = libjpeg_planes(data)
    except ImportError:
        return
    for p, r in zip(planes, ref):
        assert np.array_equal(p, r)


# ------------------------------------------------------------------------------ generated streams against libjpeg-turbo
@pytest.mark.parametrize("w,h", [(16, 16), (64, 48), (250, 130), (642, 362), (1280, 720), (1926, 1086), (3840, 2160)])
@pytest.mark.parametrize("quality,optimize", [(30, False), (75, True), (92, False), (100, True)])
def test_generated_420(uh, w, h, quality, optimize):
    pytest.importorskip("PIL.Image")
    data = encode(scene(w, h, w + quality), quality=quality, subsampling=2, optimize=optimize)
    n0 = uh.lib.uhdrb200_launch_count()
    planes = decode_host(uh.lib, data)
    passes = uh.lib.uhdrb200_launch_count() - n0 - 7
    ref = libjpeg_planes(data)
    for i, (p, r) in enumerate(zip(planes, ref)):
        bad = int((p != r).sum())
        assert bad == 0, f"plane {i}: {bad} of {p.size} samples differ ({w}x{h} q{quality})"
    print(f"{w}x{h} q{quality} optimize={optimize}: {len(data)} bytes, {passes} synchronisation passes, planes byte-identical")


@pytest.mark.parametrize("w,h", [(8, 8), (100, 60), (1283, 721), (4096, 8)])
def test_generated_grey(uh, w, h):
    pytest.importorskip("PIL.Image")
    data = encode(scene(w, h, w)[..., 1], quality=85)
    planes = decode_host(uh.lib, data)
    ref = libjpeg_planes(data)
    assert len(planes) == 1 and np.array_equal(planes[0], ref[0])


@pytest.mark.parametrize("kind", ["noise_q100", "flat", "gainmap_like"])
def test_extreme_streams(uh, kind):
    """i.i.d. noise at quality 100 (the longest codes, 16-bit ones included, hardly any EOB), a single-colour image (a
    stream of a few hundred bytes: fewer subsequences than threads) and a smooth greyscale map."""
    pytest.importorskip("PIL.Image")
    rng = np.random.Generator(np.random.PCG64(9))
    if kind == "noise_q100":
        data = encode(rng.integers(0, 256, (360, 640, 3), dtype=np.uint8), quality=100, subsampling=2, optimize=True)
    elif kind == "flat":
        data = encode(np.full((720, 1280, 3), 77, np.uint8), quality=90, subsampling=2)
    else:
        data = encode(synth.natural_map(960, 540, 3), quality=85)
    planes = decode_host(uh.lib, data)
    for p, r in zip(planes, libjpeg_planes(data)):
        assert np.array_equal(p, r)


@pytest.mark.parametrize("w,h", [(64, 48), (642, 362), (1280, 720), (3840, 2160)])
@pytest.mark.parametrize("restart", [dict(restart_marker_rows=1), dict(restart_marker_blocks=1), dict(restart_marker_blocks=7),
                                     dict(restart_marker_rows=3), dict(restart_marker_blocks=100000)])
def test_restart_intervals(uh, w, h, restart):
    """DRI streams (T.81 B.2.4.4): an interval per MCU row, per MCU, every 7 MCUs (intervals that straddle MCU rows),
    every 3 MCU rows, and an interval longer than the image (no marker is ever written). Every interval starts in a known
    state, so the decoder needs no synchronisation across them; the DC prediction restarts with each."""
    pytest.importorskip("PIL.Image")
    data = encode(scene(w, h, w + 3), quality=88, subsampling=2, **restart)
    n_rst = sum(data.count(bytes([0xFF, 0xD0 + k])) for k in range(8))
    planes = decode_host(uh.lib, data)
    for i, (p, r) in enumerate(zip(planes, libjpeg_planes(data))):
        bad = int((p != r).sum())
        assert bad == 0, f"plane {i}: {bad} of {p.size} samples differ ({w}x{h}, {restart}, ~{n_rst} markers)"


def test_restart_intervals_grey_and_q100(uh):
    pytest.importorskip("PIL.Image")
    for arr, kw in ((scene(1283, 721, 4)[..., 0], dict(quality=90, restart_marker_rows=1)),
                    (scene(640, 360, 6, noise=40.0), dict(quality=100, subsampling=2, restart_marker_blocks=3, optimize=True))):
        data = encode(arr, **kw)
        for p, r in zip(decode_host(uh.lib, data), libjpeg_planes(data)):
            assert np.array_equal(p, r)


def test_restart_marker_count_mismatch_is_an_error(uh):
    """A stream with more RSTn markers than its DRI allows is refused (libjpeg would resynchronise and warn)."""
    pytest.importorskip("PIL.Image")
    data = bytearray(encode(scene(64, 48, 8), quality=90, subsampling=2, restart_marker_rows=1))
    i = data.find(b"\xff\xdd")
    data[i + 4: i + 6] = (60000).to_bytes(2, "big")  # the interval now covers the whole image: no marker may follow
    img = abi.Image()
    out = np.zeros(64 * 48 * 3, np.uint8)
    img.data = out.ctypes.data
    buf = (C.c_ubyte * len(data)).from_buffer_copy(bytes(data))
    assert uh.lib.uhdrb200_jpeg_decode_host(buf, len(data), C.byref(img)) == abi.ERROR_DECODE


def test_many_images_one_thread_reuses_scratch(uh):
    """The calling thread's scratch buffers shrink and grow between calls: large, small, large again."""
    pytest.importorskip("PIL.Image")
    for w, h, q in [(1280, 720, 90), (32, 32, 50), (1920, 1080, 95), (64, 64, 100), (1280, 720, 90)]:
        data = encode(scene(w, h, q), quality=q, subsampling=2)
        for p, r in zip(decode_host(uh.lib, data), libjpeg_planes(data)):
            assert np.array_equal(p, r)


# ------------------------------------------------------------------------------ refusals and damaged streams
def test_refusals(uh):
    pytest.importorskip("PIL.Image")
    lib = uh.lib
    rgb = scene(64, 48, 1)
    img = abi.Image()
    out = np.zeros(64 * 48 * 3, np.uint8)
    img.data = out.ctypes.data
    for kw, code in ((dict(subsampling=0), abi.ERROR_UNSUPPORTED_FEATURE), (dict(subsampling=2, progressive=True), abi.ERROR_UNSUPPORTED_FEATURE),
                     (dict(subsampling=1), abi.ERROR_UNSUPPORTED_FEATURE)):
        data = encode(rgb, quality=90, **kw)
        buf = (C.c_ubyte * len(data)).from_buffer_copy(data)
        assert lib.uhdrb200_jpeg_decode_host(buf, len(data), C.byref(img)) == code
    junk = bytes(range(256)) * 4
    buf = (C.c_ubyte * len(junk)).from_buffer_copy(junk)
    assert lib.uhdrb200_jpeg_decode_host(buf, len(junk), C.byref(img)) == abi.ERROR_DECODE
    assert lib.uhdrb200_jpeg_decode_host(None, 4, C.byref(img)) == abi.ERROR_BAD_PTR
    img.data = None
    good = encode(rgb, quality=90, subsampling=2)
    buf = (C.c_ubyte * len(good)).from_buffer_copy(good)
    assert lib.uhdrb200_jpeg_decode_host(buf, len(good), C.byref(img)) == abi.ERROR_BAD_PTR


@pytest.mark.parametrize("damage", ["truncated", "bit_flips", "zeros"])
def test_damaged_entropy_data_stays_in_bounds(uh, damage):
    """A scan that ends early or carries garbage decodes to SOMETHING (libjpeg warns and carries on, too) without writing
    outside the planes (guard band in decode_host) or hanging."""
    pytest.importorskip("PIL.Image")
    data = bytearray(encode(scene(642, 362, 5), quality=90, subsampling=2))
    sos = data.find(b"\xff\xda")
    body = sos + 14
    if damage == "truncated":
        data = data[: body + (len(data) - body) // 3] + b"\xff\xd9"
    elif damage == "bit_flips":
        rng = np.random.Generator(np.random.PCG64(2))
        for i in rng.integers(body, len(data) - 2, 200):
            if data[i] != 0xFF and data[i - 1] != 0xFF:
                data[i] ^= 1 << int(rng.integers(0, 8))
                if data[i] == 0xFF:
                    data[i] = 0xFE
    else:
        data[body + 100: len(data) - 2] = bytes(len(data) - 2 - body - 100)
    planes = decode_host(uh.lib, bytes(data))
    assert planes[0].shape == (362, 642)


# ------------------------------------------------------------------------------ decode -> apply without leaving the device
def test_decode_dev_feeds_apply(uh, orc):
    """Primary image and gain map as JPEG bytes -> uhdrb200_jpeg_decode_dev (planes stay in device memory) ->
    uhdrb200_apply_gainmap_dev: what jpegr.cpp:693-801 does through libjpeg and host memory. Result == the oracle's
    applyGainMap on the planes libjpeg-turbo decodes."""
    pytest.importorskip("PIL.Image")
    import torch as t

    lib = uh.lib
    w, h = 1280, 720
    jp = encode(scene(w, h, 77), quality=92, subsampling=2)
    jm = encode(synth.natural_map(w // 4, h // 4, 5), quality=85)
    d_yuv = t.zeros(w * h * 3 // 2, dtype=t.uint8, device="cuda")
    d_map = t.zeros(w * h // 16, dtype=t.uint8, device="cuda")
    d_out = t.zeros(w * h, dtype=t.int32, device="cuda")
    st = t.cuda.current_stream().cuda_stream
    sdr, mp, dst = abi.Image(), abi.Image(), abi.Image()
    sdr.data, mp.data, dst.data = d_yuv.data_ptr(), d_map.data_ptr(), d_out.data_ptr()
    for data, img in ((jp, sdr), (jm, mp)):
        buf = (C.c_ubyte * len(data)).from_buffer_copy(data)
        assert lib.uhdrb200_jpeg_decode_dev(buf, len(data), C.byref(img), st) == 0, lib.uhdrb200_last_error()
    assert (sdr.width, sdr.height, sdr.chroma_stride, sdr.pixelFormat) == (w, h, w // 2, abi.PIX_FMT_YUV420)
    assert (mp.width, mp.height, mp.pixelFormat) == (w // 4, h // 4, abi.PIX_FMT_MONOCHROME)
    sdr.colorGamut = abi.COLORGAMUT_BT709
    md = abi.default_metadata(1000.0 / 203.0)
    assert lib.uhdrb200_apply_gainmap_dev(C.byref(sdr), C.byref(mp), C.byref(md), abi.OUTPUT_HDR_HLG, FLT_MAX, C.byref(dst), st) == 0
    t.cuda.synchronize()
    got = d_out.cpu().numpy().view(np.uint8)
    ref_p, ref_m = libjpeg_planes(jp), libjpeg_planes(jm)
    assert np.array_equal(d_yuv.cpu().numpy(), np.concatenate([p.ravel() for p in ref_p]))
    assert np.array_equal(d_map.cpu().numpy(), ref_m[0].ravel())
    sdr_b = np.concatenate([p.ravel() for p in ref_p])
    _, want, _ = orc.apply_gainmap(abi.host_yuv420(sdr_b, w, h, abi.COLORGAMUT_BT709), abi.host_map(ref_m[0], w // 4, h // 4), md,
                                   abi.OUTPUT_HDR_HLG, FLT_MAX)
    parity.compare_output(abi.OUTPUT_HDR_HLG, got, want, w * h, "decode_dev -> apply", min_exact=0.995)


def test_four_host_threads_decode_and_encode_concurrently(uh):
    """Every calling thread owns its context, scratch buffers and stream: four threads decoding different images (and
    encoding them again) at the same time get the results a single thread gets."""
    pytest.importorskip("PIL.Image")
    import threading

    import torch as t

    lib = uh.lib
    jobs = []
    for k in range(4):
        w, h = 640 + 64 * k, 368 + 16 * k
        data = encode(scene(w, h, 50 + k), quality=80 + 5 * k, subsampling=2)
        jobs.append((w, h, data, np.concatenate([p.ravel() for p in libjpeg_planes(data)])))
    errors = []

    def worker(k):
        try:
            t.cuda.set_device(0)
            w, h, data, want = jobs[k]
            stream = t.cuda.Stream()
            d = t.zeros(w * h * 3 // 2, dtype=t.uint8, device="cuda")
            buf = (C.c_ubyte * len(data)).from_buffer_copy(data)
            out = np.zeros(w * h * 3, np.uint8)
            n = C.c_size_t(0)
            for _ in range(6):
                img = abi.Image()
                img.data = d.data_ptr()
                d.zero_()
                t.cuda.current_stream().synchronize()
                assert lib.uhdrb200_jpeg_decode_dev(buf, len(data), C.byref(img), stream.cuda_stream) == 0, lib.uhdrb200_last_error()
                assert np.array_equal(d.cpu().numpy(), want)
                assert lib.uhdrb200_jpeg_encode_dev(C.byref(img), 77, None, 0, out.ctypes.data, out.size, C.byref(n), stream.cuda_stream) == 0
                again = bytes(out[: n.value])
                assert again[:2] == b"\xff\xd8" and again[-2:] == b"\xff\xd9"
                if k == 0:  # whole MCUs (640 x 368): the re-encoded file must be libjpeg-turbo's for the decoded planes
                    from test_gpu_jpeg_encode import pillow_encode

                    pl = [want[: w * h].reshape(h, w), want[w * h: w * h * 5 // 4].reshape(h // 2, w // 2), want[w * h * 5 // 4:].reshape(h // 2, w // 2)]
                    assert again == pillow_encode(pl, 77)
        except Exception as e:  # noqa: BLE001
            errors.append((k, repr(e)))

    ths = [threading.Thread(target=worker, args=(k,)) for k in range(4)]
    for th in ths:
        th.start()
    for th in ths:
        th.join()
    assert not errors, errors


def test_limit_size_8192(uh):
    """kMaxWidth x kMaxHeight (lib/include/ultrahdr/jpegr.h:31-32, jpegdecoderhelper.cpp:254-258): 8192 x 8192 4:2:0, the
    largest image the reference's decoder admits; 100 M samples, 1.6 M blocks, the 32-bit bit offsets and block indices at
    their largest. Encoded on the device, compared with libjpeg-turbo's file, decoded on the device, compared with
    libjpeg-turbo's planes."""
    pytest.importorskip("PIL.Image")
    from test_gpu_jpeg_encode import device_encode, pillow_encode

    w = h = 8192
    rng = np.random.Generator(np.random.PCG64(8192))
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    y = np.clip(128 + 100 * np.sin(xx / 37.0) * np.cos(yy / 29.0) + rng.normal(0, 4, (h, w)).astype(np.float32), 0, 255).astype(np.uint8)
    del yy, xx
    cy, cx = np.mgrid[0:h // 2, 0:w // 2].astype(np.float32)
    u = np.clip(128 + 60 * np.sin(cx / 41.0), 0, 255).astype(np.uint8)
    v = np.clip(128 + 60 * np.cos(cy / 33.0), 0, 255).astype(np.uint8)
    del cy, cx
    planes = [y, u, v]
    data = device_encode(uh.lib, planes, 90)
    assert data == pillow_encode(planes, 90)
    got = decode_host(uh.lib, data)
    for p, r in zip(got, libjpeg_planes(data)):
        assert np.array_equal(p, r)
    print(f"8192x8192 4:2:0 quality 90: {len(data)} bytes")
