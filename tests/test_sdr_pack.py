"""Row N3: SDR RGBA8888 pack and copy-with-restride.

CPU part: the C restatement (oracle/uhdr_oracle.c) against the reference's own routine
UltraHdrAppInput::convertYuv420ToRGBImage (examples/ultrahdr_app.cpp:682-734) compiled unmodified
into oracle/_ref/libuhdr_ref_app.so, plus hand-computed known answers.
GPU part: the CUDA path through the C ABI, bit-exact against the restatement."""
import ctypes as C

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth
from oracle import pyoracle


def _planes(buf, w, h, ls, cs):
    y = buf[: ls * h]
    u = buf[ls * h: ls * h + cs * (h // 2)]
    v = buf[ls * h + cs * (h // 2):]
    return y, u, v


@pytest.mark.parametrize("w,h,ls,cs", [(64, 32, 64, 32), (70, 38, 75, 39), (2, 2, 2, 1), (256, 130, 260, 131)])
def test_oracle_equals_reference_app_routine(orc, w, h, ls, cs):
    if not pyoracle.have_ref_app():
        pytest.skip("oracle/_ref/libuhdr_ref_app.so not built (reference tree absent)")
    buf = np.random.Generator(np.random.PCG64(7)).integers(0, 256, ls * h + cs * h, dtype=np.uint8)
    y, u, v = _planes(buf, w, h, ls, cs)
    want = pyoracle.ref_app_yuv420_to_rgba8888(y, u, v, w, h, ls, cs, cs)
    st, got, dest = orc.yuv420_to_rgba8888(abi.host_yuv420(buf, w, h, 1, ls, cs))
    assert st == 0 and (dest.width, dest.height, dest.colorGamut) == (w, h, 1)
    assert np.array_equal(got, want)


def test_oracle_known_answers(orc):
    """Grey stays grey, pure chroma excursions clip: values worked out by hand from the 601 matrix."""
    w = h = 2
    for yv, uv, vv, rgb in [(0, 128, 128, (0, 0, 0)), (255, 128, 128, (255, 255, 255)), (128, 128, 128, (128, 128, 128)),
                            (128, 128, 255, (255, 37, 128)), (128, 255, 128, (128, 84, 255)), (16, 0, 0, (0, 151, 0))]:
        buf = np.array([yv] * 4 + [uv] + [vv], dtype=np.uint8)
        st, got, _ = orc.yuv420_to_rgba8888(abi.host_yuv420(buf, w, h, 0))
        px = int(got[0, 0])
        assert st == 0 and (px & 0xff, (px >> 8) & 0xff, (px >> 16) & 0xff, px >> 24) == rgb + (255,), (yv, uv, vv, hex(px))
        assert (got == px).all()


def test_oracle_copy_yuv420(orc):
    w, h, ls, cs, dls, dcs = 70, 38, 75, 39, 128, 64
    src_b = np.random.Generator(np.random.PCG64(9)).integers(1, 256, ls * h + cs * h, dtype=np.uint8)
    dst_b = np.full(dls * h + dcs * h, 0xAA, dtype=np.uint8)
    src, dst = abi.host_yuv420(src_b, w, h, 2, ls, cs), abi.host_yuv420(dst_b, w, h, 0, dls, dcs)
    assert orc.copy_yuv420(src, dst) == 0 and dst.colorGamut == 2
    sy, su, sv = _planes(src_b, w, h, ls, cs)
    dy, du, dv = _planes(dst_b, w, h, dls, dcs)
    assert np.array_equal(dy.reshape(h, dls)[:, :w], sy.reshape(h, ls)[:, :w]) and not dy.reshape(h, dls)[:, w:].any()
    for d, s in ((du, su), (dv, sv)):
        assert np.array_equal(d.reshape(h // 2, dcs)[:, : w // 2], s.reshape(h // 2, cs)[:, : w // 2])
        assert not d.reshape(h // 2, dcs)[:, w // 2:].any()
    bad = abi.host_yuv420(dst_b, w, h, 0, 64, 32)
    assert orc.copy_yuv420(src, bad) == abi.ERROR_INVALID_STRIDE


# ------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("w,h,ls,cs,os_", [(256, 128, 256, 128, 0), (64, 32, 68, 34, 72), (70, 38, 75, 39, 0), (2, 2, 2, 1, 0),
                                           (1920, 1080, 1920, 960, 0)])
@pytest.mark.parametrize("kind", ["natural", "uniform"])
def test_gpu_yuv420_to_rgba8888_exact(uh, orc, w, h, ls, cs, os_, kind):
    if kind == "natural" and ls == w and cs == w // 2 and w >= 64:
        buf, _ = synth.natural_pair(w, h, 61)
    else:
        buf = np.random.Generator(np.random.PCG64(62)).integers(0, 256, ls * h + cs * h, dtype=np.uint8)
    img = abi.host_yuv420(buf, w, h, 0, ls, cs)
    st, want, _ = orc.yuv420_to_rgba8888(img, os_)
    stride = max(os_, w)
    out = np.full(h * stride, 0xDEADBEEF, dtype=np.uint32)
    dest = abi.Image()
    dest.data, dest.luma_stride = out.ctypes.data, os_
    assert uh.yuv420ToRgba8888(img, dest) == 0 and st == 0
    assert (dest.width, dest.height, dest.colorGamut) == (w, h, 0)
    got = out.reshape(h, stride)
    assert np.array_equal(got[:, :w], want[:, :w])
    assert (got[:, w:] == 0xDEADBEEF).all()  # bytes beyond width are the caller's


@pytest.mark.gpu
def test_gpu_copy_yuv420_exact(uh, orc):
    for (w, h, ls, cs, dls, dcs) in [(70, 38, 75, 39, 128, 64), (256, 128, 256, 128, 256, 128), (1920, 1080, 1920, 960, 1920, 960)]:
        src_b = np.random.Generator(np.random.PCG64(63)).integers(1, 256, ls * h + cs * h, dtype=np.uint8)
        want_b = np.full(dls * h + dcs * h, 0xAA, dtype=np.uint8)
        got_b = want_b.copy()
        src = abi.host_yuv420(src_b, w, h, 1, ls, cs)
        assert orc.copy_yuv420(src, abi.host_yuv420(want_b, w, h, 0, dls, dcs)) == 0
        dst = abi.host_yuv420(got_b, w, h, 0, dls, dcs)
        assert uh.copyYuv420(src, dst) == 0 and dst.colorGamut == 1
        assert np.array_equal(got_b, want_b)


@pytest.mark.gpu
def test_gpu_sdr_pack_argument_checks(uh):
    buf = np.zeros(64 * 32 * 3 // 2, dtype=np.uint8)
    img = abi.host_yuv420(buf, 64, 32, 0)
    out = np.zeros(64 * 32, dtype=np.uint32)
    dest = abi.Image()
    assert uh.yuv420ToRgba8888(img, dest) == abi.ERROR_BAD_PTR
    dest.data = out.ctypes.data
    assert uh.yuv420ToRgba8888(None, dest) == abi.ERROR_BAD_PTR
    odd = abi.host_yuv420(buf, 63, 32, 0, 64, 32)
    assert uh.yuv420ToRgba8888(odd, dest) == abi.ERROR_UNSUPPORTED_WIDTH_HEIGHT
