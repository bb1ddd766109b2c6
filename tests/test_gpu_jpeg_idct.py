"""Row N4 on the GPU: uhdrb200_jpeg_idct_{host,dev} (dequantise + ISLOW 8x8 IDCT + range-limit into planar samples)
against the oracle — which tests/test_jpeg_idct.py pins bit for bit to libjpeg-turbo's own output — on the reference's
JPEG fixtures, on random coefficient planes with ragged sizes and strides, and end to end: coefficients -> planes ->
applyGainMap without the decoded image ever leaving the device."""
import ctypes as C

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, refdata, synth

import jpeg_coeffs
import parity
from test_jpeg_idct import need_fixtures, oracle_planes

pytestmark = pytest.mark.gpu
FLT_MAX = 3.4028234663852886e38


def host_idct(lib, comps, strides=None, fill=0x5A):
    n = len(comps)
    arr = (abi.JpegComponent * n)()
    keep, planes = [], []
    for i, c in enumerate(comps):
        co, q = np.ascontiguousarray(c.coeffs), np.ascontiguousarray(c.quant)
        stride = (strides[i] if strides else None) or c.width
        plane = np.full((c.height, stride), fill, np.uint8)
        keep += [co, q]
        planes.append(plane)
        arr[i].coeffs, arr[i].quant = co.ctypes.data, q.ctypes.data
        arr[i].blocks_w, arr[i].blocks_h = co.shape[1], co.shape[0]
        arr[i].plane, arr[i].width, arr[i].height, arr[i].stride = plane.ctypes.data, c.width, c.height, stride
    st = lib.uhdrb200_jpeg_idct_host(n, arr)
    assert st == 0, lib.uhdrb200_last_error()
    return planes


@need_fixtures
@pytest.mark.parametrize("name", refdata.JPEGS)
def test_fixture_planes_bit_exact(uh, orc, name):
    w, h, comps = jpeg_coeffs.decode(open(refdata.jpeg_path(name), "rb").read())
    want = oracle_planes(orc, comps)
    got = host_idct(uh.lib, comps)
    for g, wv in zip(got, want):
        assert np.array_equal(g, wv)


class Rnd:
    pass


@pytest.mark.parametrize("w,h,pad", [(8, 8, 0), (17, 9, 3), (100, 60, 0), (250, 130, 6), (1920, 1080, 0), (33, 1, 5)])
def test_random_coefficients_ragged(uh, orc, w, h, pad):
    """Random sparse coefficients incl. values that leave the 8-bit range (libjpeg's range-limit wrap), sizes that are not
    multiples of 8, padded strides: bit exact, and bytes outside the visible samples untouched."""
    rng = np.random.Generator(np.random.PCG64(w * 1000 + h))
    comps = []
    for k in range(3):
        c = Rnd()
        cw, ch = (w, h) if k == 0 else ((w + 1) // 2, (h + 1) // 2)
        bw, bh = -(-cw // 8) + (k == 0), -(-ch // 8)  # luma: one spare block column, like MCU padding
        co = (rng.integers(-300, 300, (bh, bw, 64)) * (rng.random((bh, bw, 64)) < 0.15)).astype(np.int16)
        co[..., 0] = rng.integers(-1024, 1024, (bh, bw))
        c.coeffs, c.quant = co, rng.integers(1, 64, 64).astype(np.uint16)
        c.width, c.height = cw, ch
        comps.append(c)
    want = oracle_planes(orc, comps)
    strides = [c.width + pad for c in comps]
    got = host_idct(uh.lib, comps, strides)
    for c, g, wv in zip(comps, got, want):
        assert np.array_equal(g[:, : c.width], wv)
        assert (g[:, c.width:] == 0x5A).all()


@pytest.mark.parametrize("wide", [(True, False, False), (False, False, True), (True, True, True)])
def test_sixteen_bit_quant_tables(uh, orc, wide):
    """Quantisation steps above 255 (a 16-bit DQT, Pq = 1) in some or all components: the kernel's plain-multiply
    dequantisation instead of its packed one; small coefficients keep the sums inside libjpeg's defined range."""
    rng = np.random.Generator(np.random.PCG64(sum(wide) * 7 + wide[0]))
    comps = []
    for k in range(3):
        c = Rnd()
        cw, ch = (200, 120) if k == 0 else (100, 60)
        bw, bh = -(-cw // 8), -(-ch // 8)
        hi = 4000 if wide[k] else 256
        q = rng.integers(1, hi, 64).astype(np.uint16)
        if wide[k]:
            q[rng.integers(0, 64)] = 65535
        else:
            q[rng.integers(0, 64)] = 255
        co = (rng.integers(-3, 4, (bh, bw, 64)) * (rng.random((bh, bw, 64)) < 0.1)).astype(np.int16)
        co[..., 0] = rng.integers(-1024, 1024, (bh, bw)) // np.maximum(1, q[0] // 8)
        co[q[None, None, :].repeat(bh, 0).repeat(bw, 1) > 2000] = 0
        c.coeffs, c.quant, c.width, c.height = co, q, cw, ch
        comps.append(c)
    want = oracle_planes(orc, comps)
    got = host_idct(uh.lib, comps, [c.width for c in comps])
    for g, wv in zip(got, want):
        assert np.array_equal(g, wv)
    assert len(np.unique(want[0])) > 16


@pytest.mark.parametrize("co_off,pl_off", [(1, 0), (3, 1), (0, 3)])
def test_device_pointers_of_any_alignment(uh, orc, co_off, pl_off):
    """Device-pointer callers: coefficient arrays that are only 2-byte aligned (the kernel's asynchronous 16-byte copies
    need 16) and planes at odd addresses (byte stores instead of 8-byte ones): same samples."""
    import torch as t

    rng = np.random.Generator(np.random.PCG64(co_off * 10 + pl_off))
    comps, arr, keep, outs = [], (abi.JpegComponent * 3)(), [], []
    for k in range(3):
        c = Rnd()
        cw, ch = (333, 70) if k == 0 else (167, 35)
        bw, bh = -(-cw // 8), -(-ch // 8)
        co = (rng.integers(-200, 200, (bh, bw, 64)) * (rng.random((bh, bw, 64)) < 0.2)).astype(np.int16)
        co[..., 0] = rng.integers(-900, 900, (bh, bw))
        c.coeffs, c.quant, c.width, c.height = co, rng.integers(1, 40, 64).astype(np.uint16), cw, ch
        comps.append(c)
        dco = t.zeros(co.size + 8, dtype=t.int16, device="cuda")
        dco[co_off: co_off + co.size] = t.from_numpy(co.ravel()).cuda()
        dq = t.from_numpy(c.quant.astype(np.int16)).cuda()
        dpl = t.full((cw * ch + 8,), 0x5A, dtype=t.uint8, device="cuda")
        keep += [dco, dq]
        outs.append(dpl)
        arr[k].coeffs, arr[k].quant = dco.data_ptr() + 2 * co_off, dq.data_ptr()
        arr[k].blocks_w, arr[k].blocks_h = bw, bh
        arr[k].plane, arr[k].width, arr[k].height, arr[k].stride = dpl.data_ptr() + pl_off, cw, ch, cw
    assert uh.lib.uhdrb200_jpeg_idct_dev(3, arr, t.cuda.current_stream().cuda_stream) == 0, uh.lib.uhdrb200_last_error()
    t.cuda.synchronize()
    for c, d, wv in zip(comps, outs, oracle_planes(orc, comps)):
        g = d.cpu().numpy()
        assert np.array_equal(g[pl_off: pl_off + c.width * c.height].reshape(c.height, c.width), wv)
        assert (g[:pl_off] == 0x5A).all() and (g[pl_off + c.width * c.height:] == 0x5A).all()


def test_error_codes(uh):
    lib = uh.lib
    assert lib.uhdrb200_jpeg_idct_host(0, None) == abi.NO_ERROR
    assert lib.uhdrb200_jpeg_idct_host(1, None) == abi.ERROR_BAD_PTR
    arr = (abi.JpegComponent * 4)()
    assert lib.uhdrb200_jpeg_idct_host(4, arr) == abi.ERROR_UNSUPPORTED_FEATURE  # host path: up to 3 components
    co, q, pl = np.zeros((1, 1, 64), np.int16), np.ones(64, np.uint16), np.zeros((8, 8), np.uint8)
    arr[0].coeffs, arr[0].quant, arr[0].plane = co.ctypes.data, q.ctypes.data, pl.ctypes.data
    arr[0].blocks_w = arr[0].blocks_h = 1
    arr[0].width, arr[0].height, arr[0].stride = 9, 8, 9
    assert lib.uhdrb200_jpeg_idct_host(1, arr) == abi.ERROR_RESOLUTION_MISMATCH
    arr[0].width, arr[0].stride = 8, 4
    assert lib.uhdrb200_jpeg_idct_host(1, arr) == abi.ERROR_INVALID_STRIDE
    arr[0].stride = 8
    assert lib.uhdrb200_jpeg_idct_host(1, arr) == abi.NO_ERROR and (pl == 128).all()


@need_fixtures
def test_decode_to_apply_stays_on_device(uh, orc):
    """jpeg_image.jpg (1280x720, 4:2:0): coefficients -> Y | U | V planes on the device (jpeg_idct_dev) -> applyGainMap on
    the same device buffers (what jpegr.cpp:693-801 does through host memory). Result == the oracle's applyGainMap on the
    planes libjpeg-turbo decodes."""
    import torch as t

    lib = uh.lib
    w, h, comps = jpeg_coeffs.decode(open(refdata.jpeg_path("jpeg_image.jpg"), "rb").read())
    d_yuv = t.zeros(w * h * 3 // 2, dtype=t.uint8, device="cuda")
    offs, arr, keep = [0, w * h, w * h * 5 // 4], (abi.JpegComponent * 3)(), []
    for i, c in enumerate(comps):
        dc = t.from_numpy(np.ascontiguousarray(c.coeffs)).cuda()
        dq = t.from_numpy(c.quant.astype(np.int16)).cuda()
        keep += [dc, dq]
        arr[i].coeffs, arr[i].quant = dc.data_ptr(), dq.data_ptr()
        arr[i].blocks_w, arr[i].blocks_h = c.coeffs.shape[1], c.coeffs.shape[0]
        arr[i].plane, arr[i].width, arr[i].height, arr[i].stride = d_yuv.data_ptr() + offs[i], c.width, c.height, c.width
    st = t.cuda.current_stream().cuda_stream
    assert lib.uhdrb200_jpeg_idct_dev(3, arr, st) == 0, lib.uhdrb200_last_error()
    mp_b = synth.natural_map(w // 4, h // 4, 5)
    d_map = t.from_numpy(mp_b.ravel()).cuda()
    d_out = t.zeros(w * h, dtype=t.int32, device="cuda")
    sdr, mp, dst = abi.Image(), abi.Image(), abi.Image()
    sdr.data, sdr.chroma_data, sdr.width, sdr.height, sdr.colorGamut = d_yuv.data_ptr(), d_yuv.data_ptr() + w * h, w, h, 0
    sdr.luma_stride, sdr.chroma_stride, sdr.pixelFormat = w, w // 2, abi.PIX_FMT_YUV420
    mp.data, mp.width, mp.height, mp.luma_stride, mp.pixelFormat = d_map.data_ptr(), w // 4, h // 4, w // 4, abi.PIX_FMT_MONOCHROME
    dst.data = d_out.data_ptr()
    md = abi.default_metadata(1000.0 / 203.0)
    assert uh.applyGainMapDevice(sdr, mp, md, abi.OUTPUT_HDR_HLG, FLT_MAX, dst, st) == 0
    t.cuda.synchronize()
    planes = oracle_planes(orc, comps)
    yuv = np.concatenate([p.ravel() for p in planes])
    assert np.array_equal(d_yuv.cpu().numpy(), yuv)
    _, want, _ = orc.apply_gainmap(abi.host_yuv420(yuv, w, h, 0), abi.host_map(mp_b, w // 4, h // 4), md, abi.OUTPUT_HDR_HLG, FLT_MAX)
    exact = parity.compare_output(abi.OUTPUT_HDR_HLG, d_out.cpu().numpy().view(np.uint8), want, w * h, "decode -> apply on device", min_exact=0.995)
    print(f"jpeg_image.jpg decode -> apply on device: exact {exact:.5f}")
