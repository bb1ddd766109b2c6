"""GPU: no kernel writes outside its output planes. Every output (and every in-place plane) sits between two 4 KiB guard
bands filled with 0xA5 inside one device allocation; after the call the bands must be untouched. Ragged sizes (last tile
row / column partly empty), every apply kernel family (scale-4 q-domain, any-scale specialisations 1 / 2 and the
run-time scale, analytic) and the other stages. compute-sanitizer is not available on the pool, this is its stand-in
for writes."""
import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth

pytestmark = pytest.mark.gpu
FLT_MAX = 3.4028234663852886e38
G = 4096


@pytest.fixture(scope="module")
def t():
    import torch

    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    return torch


class Guarded:
    """payload of n bytes between two guard bands; .ptr is 4 KiB-aligned relative to the allocation"""

    def __init__(self, t, n, fill=None):
        self.t, self.n = t, n
        self.buf = t.full((n + 2 * G,), 0xA5, dtype=t.uint8, device="cuda")
        if fill is not None:
            self.buf[G:G + n] = t.from_numpy(np.ascontiguousarray(fill).view(np.uint8).ravel()).cuda()
        self.ptr = self.buf.data_ptr() + G

    def check(self, what):
        self.t.cuda.synchronize()
        lo, hi = self.buf[:G].cpu().numpy(), self.buf[G + self.n:].cpu().numpy()
        assert (lo == 0xA5).all(), f"{what}: wrote below the plane"
        assert (hi == 0xA5).all(), f"{what}: wrote above the plane"

    def payload(self):
        return self.buf[G:G + self.n].cpu().numpy()


def _sdr_img(g, w, h, gamut=0):
    img = abi.Image()
    img.data, img.chroma_data = g.ptr, g.ptr + w * h
    img.width, img.height, img.colorGamut = w, h, gamut
    img.luma_stride, img.chroma_stride, img.pixelFormat = w, w // 2, abi.PIX_FMT_YUV420
    return img


@pytest.mark.parametrize("w,h", [(132, 68), (260, 12), (516, 36), (8, 8)])
@pytest.mark.parametrize("scale", [1, 2, 4, 3])
def test_apply_writes_stay_inside(uh, t, w, h, scale):
    if scale == 3:
        w, h = w // 4 * 12, h // 4 * 12  # multiples of 4 and of 3: the run-time-scale kernel
    sdr_b, _ = synth.uniform_pair(w, h, 5)
    g_sdr = Guarded(t, sdr_b.size, sdr_b)
    mw, mh = w // scale, h // scale
    mp = np.random.Generator(np.random.PCG64(6)).integers(0, 256, mw * mh, dtype=np.uint8)
    g_map = Guarded(t, mp.size, mp)
    m_img = abi.Image()
    m_img.data, m_img.width, m_img.height, m_img.luma_stride = g_map.ptr, mw, mh, mw
    m_img.pixelFormat = abi.PIX_FMT_MONOCHROME
    md = abi.default_metadata(4.926108)
    st = t.cuda.current_stream().cuda_stream
    for fmt in (abi.OUTPUT_HDR_LINEAR, abi.OUTPUT_HDR_PQ, abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR_RGB_10BIT):
        g_out = Guarded(t, w * h * abi.out_bytes_per_pixel(fmt))
        o = abi.Image()
        o.data = g_out.ptr
        assert uh.applyGainMapDevice(_sdr_img(g_sdr, w, h), m_img, md, fmt, FLT_MAX, o, st) == 0
        g_out.check(f"apply fmt={fmt} scale={scale} {w}x{h}")
        assert (g_out.payload() != 0xA5).any()
    g_sdr.check("apply: SDR input")
    g_map.check("apply: map input")


@pytest.mark.parametrize("w,h", [(132, 68), (260, 12), (516, 36), (8, 8)])
def test_other_stages_write_inside(uh, t, w, h):
    sdr_b, hdr_b = synth.uniform_pair(w, h, 9)
    st = t.cuda.current_stream().cuda_stream
    g_sdr, g_hdr = Guarded(t, sdr_b.size, sdr_b), Guarded(t, hdr_b.size * 2, hdr_b)
    h_img = abi.Image()
    h_img.data, h_img.chroma_data = g_hdr.ptr, g_hdr.ptr + 2 * w * h
    h_img.width, h_img.height, h_img.colorGamut = w, h, abi.COLORGAMUT_BT2100
    h_img.luma_stride, h_img.chroma_stride, h_img.pixelFormat = w, w, abi.PIX_FMT_P010
    # generate
    g_map = Guarded(t, w * h // 16)
    md, m_img = abi.Metadata(), abi.Image()
    m_img.data = g_map.ptr
    assert uh.generateGainMapDevice(_sdr_img(g_sdr, w, h), h_img, abi.TF_HLG, md, m_img, False, st) == 0
    g_map.check("generate")
    # toneMap and the fused API-0 pass (with and without convertYuv)
    for fused, to601 in ((False, False), (True, False), (True, True)):
        g_tm, g_m2 = Guarded(t, sdr_b.size), Guarded(t, w * h // 16)
        tm = _sdr_img(g_tm, w, h, -1)
        if fused:
            md2, m2 = abi.Metadata(), abi.Image()
            m2.data = g_m2.ptr
            assert uh.toneMapGenerateGainMapDevice(h_img, abi.TF_HLG, tm, md2, m2, st, to601) == 0
            g_m2.check(f"fused map 601={to601}")
        else:
            assert uh.toneMapDevice(h_img, tm, st) == 0
        g_tm.check(f"toneMap fused={fused} 601={to601}")
    # convertYuv in place, SDR pack
    assert uh.convertYuvDevice(_sdr_img(g_sdr, w, h), abi.COLORGAMUT_BT709, abi.COLORGAMUT_P3, st) == 0
    g_sdr.check("convertYuv in place")
    g_rgba = Guarded(t, w * h * 4)
    o = abi.Image()
    o.data = g_rgba.ptr
    assert uh.yuv420ToRgba8888Device(_sdr_img(g_sdr, w, h), o, st) == 0
    g_rgba.check("SDR pack")
    g_hdr.check("P010 input")


@pytest.mark.parametrize("w,h", [(64, 8), (128, 36), (192, 68), (1024, 20), (960, 540)])
def test_pipeline_variant_writes_stay_inside(uh, t, w, h):
    """The TMA-fed variant of the scale-4 kernel (test hook Q_PIPE; layouts with width % 64 == 0): outputs between guard
    bands, partial strips (16 of 32 cells), short walks, and byte-identical to the default kernel."""
    sdr_b, _ = synth.natural_pair(w, h, 15)
    g_sdr = Guarded(t, sdr_b.size, sdr_b)
    mp = synth.natural_map(w // 4, h // 4, 16)
    g_map = Guarded(t, mp.size, mp)
    m_img = abi.Image()
    m_img.data, m_img.width, m_img.height, m_img.luma_stride = g_map.ptr, w // 4, h // 4, w // 4
    m_img.pixelFormat = abi.PIX_FMT_MONOCHROME
    md = abi.default_metadata(4.926108)
    st = t.cuda.current_stream().cuda_stream
    for fmt in (abi.OUTPUT_HDR_LINEAR, abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR_RGB_10BIT):
        outs = []
        for mode in (abi.APPLY_Q_PIPE, abi.APPLY_AUTO):
            assert uh.lib.uhdrb200_test_set_apply_mode(mode) == 0
            g_out = Guarded(t, w * h * abi.out_bytes_per_pixel(fmt))
            o = abi.Image()
            o.data = g_out.ptr
            assert uh.applyGainMapDevice(_sdr_img(g_sdr, w, h), m_img, md, fmt, FLT_MAX, o, st) == 0
            g_out.check(f"apply mode={mode} fmt={fmt} {w}x{h}")
            outs.append(g_out.payload())
        assert uh.lib.uhdrb200_test_set_apply_mode(abi.APPLY_AUTO) == 0
        assert np.array_equal(outs[0], outs[1]), f"pipeline variant differs from the default kernel, fmt={fmt} {w}x{h}"


@pytest.mark.parametrize("w,h", [(8, 8), (17, 9), (250, 130), (264, 7)])
def test_jpeg_idct_writes_stay_inside(uh, t, w, h):
    """Row N4 through the device entry point: ragged widths / heights inside whole-MCU coefficient arrays, planes between
    guard bands, padded stride bytes untouched."""
    rng = np.random.Generator(np.random.PCG64(w + h))
    bw, bh = -(-w // 16) * 2, -(-h // 16) * 2  # whole 16x16 MCUs
    co = (rng.integers(-200, 200, (bh, bw, 64)) * (rng.random((bh, bw, 64)) < 0.2)).astype(np.int16)
    q = rng.integers(1, 40, 64).astype(np.uint16)
    stride = w + 5
    n = stride * (h - 1) + w
    g_plane = Guarded(t, n, np.full(n, 0x5A, np.uint8))
    d_co, d_q = t.from_numpy(co).cuda(), t.from_numpy(q.astype(np.int16)).cuda()
    c = (abi.JpegComponent * 1)()
    c[0].coeffs, c[0].quant, c[0].blocks_w, c[0].blocks_h = d_co.data_ptr(), d_q.data_ptr(), bw, bh
    c[0].plane, c[0].width, c[0].height, c[0].stride = g_plane.ptr, w, h, stride
    assert uh.lib.uhdrb200_jpeg_idct_dev(1, c, t.cuda.current_stream().cuda_stream) == 0
    g_plane.check(f"jpeg idct {w}x{h}")
    got = np.concatenate([g_plane.payload(), np.full(stride - w, 0x5A, np.uint8)]).reshape(h, stride)
    assert (got[:, w:] == 0x5A).all() and (got[:, :w] != 0x5A).any()
