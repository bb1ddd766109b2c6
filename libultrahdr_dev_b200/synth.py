"""Deterministic synthetic inputs for tests and bench (no datasets exist offline).

Two generators:
  * ``natural``: smooth scene-linear HDR + a smooth gain field, so the encoded gain map spans the
    whole 0..255 range instead of saturating; SDR = clip(HDR * peak/203 / gain).
  * ``uniform``: every sample i.i.d. over the full code range (P010 10-bit 0..1023 incl. values
    outside the legal 64..940 range; 420 0..255) — exercises the clamps and the un-clamped
    getP010Pixel (gainmapmath.cpp:593-600).
Seed convention: 0xB200 + image_index, NumPy PCG64.
"""
import numpy as np

from . import abi

_YW = {0: (0.2126, 0.7152, 0.0722), 1: (0.299, 0.587, 0.114), 2: (0.2627, 0.6780, 0.0593)}
_CB = {0: 1.8556, 1: 1.772, 2: 1.8814}
_CR = {0: 1.5748, 1: 1.402, 2: 1.4746}


def _rng(index):
    return np.random.Generator(np.random.PCG64(0xB200 + int(index)))


def _srgb_oetf(x):
    x = np.clip(x, 0.0, 1.0)
    return np.where(x <= 0.0031308, 12.92 * x, 1.055 * np.power(x, 1 / 2.4) - 0.055)


def _hlg_oetf(x):
    x = np.clip(x, 0.0, 1.0)
    a, b, c = 0.17883277, 0.28466892, 0.55991073
    return np.where(x <= 1 / 12, np.sqrt(3 * x), a * np.log(np.maximum(12 * x - b, 1e-9)) + c)


def _pq_oetf(x):
    x = np.clip(x, 0.0, 1.0)
    m1, m2 = 2610 / 16384, 2523 / 4096 * 128
    c1, c2, c3 = 3424 / 4096, 2413 / 4096 * 32, 2392 / 4096 * 32
    p = np.power(x, m1)
    return np.power((c1 + c2 * p) / (1 + c3 * p), m2)


def _smooth_field(rng, h, w, cells=6):
    """Low-frequency field in [0,1]: bilinear upsample of a small random grid."""
    gy, gx = cells + 1, cells + 1
    grid = rng.random((gy, gx)).astype(np.float32)
    ys = np.linspace(0, cells, h, dtype=np.float32)
    xs = np.linspace(0, cells, w, dtype=np.float32)
    y0 = np.minimum(ys.astype(np.int32), cells - 1)
    x0 = np.minimum(xs.astype(np.int32), cells - 1)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    g00 = grid[y0][:, x0]
    g01 = grid[y0][:, x0 + 1]
    g10 = grid[y0 + 1][:, x0]
    g11 = grid[y0 + 1][:, x0 + 1]
    return (g00 * (1 - fy) * (1 - fx) + g01 * (1 - fy) * fx + g10 * fy * (1 - fx) + g11 * fy * fx).astype(np.float32)


def _to_yuv420_planes(rgb_gamma, matrix):
    """rgb_gamma: (3,h,w) in [0,1] -> y (h,w), u,v (h/2,w/2) floats; y in [0,1], u,v in [-.5,.5]."""
    yr, yg, yb = _YW[matrix]
    r, g, b = rgb_gamma
    y = yr * r + yg * g + yb * b
    u = (b - y) / _CB[matrix]
    v = (r - y) / _CR[matrix]
    h, w = y.shape
    u = u.reshape(h // 2, 2, w // 2, 2).mean(axis=(1, 3))
    v = v.reshape(h // 2, 2, w // 2, 2).mean(axis=(1, 3))
    return y, u, v


def pack_yuv420(y, u, v, luma_stride=None, chroma_stride=None):
    """8-bit full-range planar buffer Y | U | V with the given strides (padding bytes = 0xAA)."""
    h, w = y.shape
    ls = luma_stride or w
    cs = chroma_stride or ls // 2
    buf = np.full(abi.yuv420_size(w, h, ls, cs), 0xAA, dtype=np.uint8)
    buf[: ls * h].reshape(h, ls)[:, :w] = y
    off = ls * h
    buf[off: off + cs * (h // 2)].reshape(h // 2, cs)[:, : w // 2] = u
    off += cs * (h // 2)
    buf[off: off + cs * (h // 2)].reshape(h // 2, cs)[:, : w // 2] = v
    return buf


def pack_p010(y10, u10, v10, luma_stride=None, chroma_stride=None, low_bits=None):
    """P010 buffer Y | interleaved UV, 10 MSBs of each u16 (padding = 0xAAAA)."""
    h, w = y10.shape
    ls = luma_stride or w
    cs = chroma_stride or ls
    buf = np.full(abi.p010_size(w, h, ls, cs), 0xAAAA, dtype=np.uint16)
    buf[: ls * h].reshape(h, ls)[:, :w] = y10.astype(np.uint16) << 6
    uv = np.empty((h // 2, w), dtype=np.uint16)
    uv[:, 0::2] = u10.astype(np.uint16) << 6
    uv[:, 1::2] = v10.astype(np.uint16) << 6
    buf[ls * h: ls * h + cs * (h // 2)].reshape(h // 2, cs)[:, :w] = uv
    if low_bits is not None:  # garbage in the 6 LSBs must be ignored by >>6
        buf |= low_bits.astype(np.uint16) & 0x3F
    return buf


def uniform_pair(width, height, index=0, sdr_luma_stride=None, sdr_chroma_stride=None,
                 hdr_luma_stride=None, hdr_chroma_stride=None, dirty_low_bits=False):
    """-> (yuv420 u8 buffer, p010 u16 buffer), every sample uniform over its full code range."""
    rng = _rng(index)
    h, w = height, width
    y = rng.integers(0, 256, (h, w), dtype=np.uint8)
    u = rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8)
    v = rng.integers(0, 256, (h // 2, w // 2), dtype=np.uint8)
    sdr = pack_yuv420(y, u, v, sdr_luma_stride, sdr_chroma_stride)
    y10 = rng.integers(0, 1024, (h, w), dtype=np.uint16)
    u10 = rng.integers(0, 1024, (h // 2, w // 2), dtype=np.uint16)
    v10 = rng.integers(0, 1024, (h // 2, w // 2), dtype=np.uint16)
    hdr = pack_p010(y10, u10, v10, hdr_luma_stride, hdr_chroma_stride)
    if dirty_low_bits:
        hdr |= rng.integers(0, 64, hdr.shape, dtype=np.uint16)
    return sdr, hdr


def natural_pair(width, height, index=0, hdr_tf=abi.TF_HLG, sdr_gamut=abi.COLORGAMUT_BT709,
                 hdr_gamut=abi.COLORGAMUT_BT2100, sdr_luma_stride=None, sdr_chroma_stride=None,
                 hdr_luma_stride=None, hdr_chroma_stride=None):
    """-> (yuv420 u8 buffer, p010 u16 buffer) of a smooth HDR scene and its tone-mapped SDR twin.

    The gamut arguments only select the YUV matrices used for encoding (colour primaries are not
    converted: the point is code coverage of the arithmetic, not colourimetry).
    """
    rng = _rng(index)
    h, w = height, width
    peak = 10000.0 if hdr_tf == abi.TF_PQ else 1000.0
    max_boost = peak / 203.0
    base = _smooth_field(rng, h, w, 5)
    noise = (rng.random((h, w), dtype=np.float32) - 0.5) * 0.1
    lum = np.clip(base * base * (1.0 + noise), 0.0, 1.0)  # scene-linear, relative to peak
    hl = rng.random((h, w), dtype=np.float32) > 0.9995  # sparse highlights
    lum = np.where(hl, 1.0, lum).astype(np.float32)
    tint = np.stack([0.7 + 0.3 * _smooth_field(rng, h, w, 3) for _ in range(3)])
    hdr_lin = np.clip(lum[None] * tint, 0.0, 1.0).astype(np.float32)
    gain = np.exp(np.log(max_boost) * _smooth_field(rng, h, w, 4)).astype(np.float32)  # [1, max_boost]
    sdr_lin = np.clip(hdr_lin * (peak / 203.0) / gain[None], 0.0, 1.0)

    if hdr_tf == abi.TF_HLG:
        hdr_gamma = _hlg_oetf(hdr_lin)
    elif hdr_tf == abi.TF_PQ:
        hdr_gamma = _pq_oetf(hdr_lin)
    else:
        hdr_gamma = hdr_lin
    hy, hu, hv = _to_yuv420_planes(hdr_gamma, hdr_gamut)
    y10 = np.clip(np.rint(64 + 876 * hy), 0, 1023)
    u10 = np.clip(np.rint(64 + 896 * (hu + 0.5)), 0, 1023)
    v10 = np.clip(np.rint(64 + 896 * (hv + 0.5)), 0, 1023)
    hdr = pack_p010(y10, u10, v10, hdr_luma_stride, hdr_chroma_stride)

    sy, su, sv = _to_yuv420_planes(_srgb_oetf(sdr_lin), sdr_gamut)
    y8 = np.clip(np.rint(255 * sy), 0, 255).astype(np.uint8)
    u8 = np.clip(np.rint(255 * su + 128), 0, 255).astype(np.uint8)
    v8 = np.clip(np.rint(255 * sv + 128), 0, 255).astype(np.uint8)
    sdr = pack_yuv420(y8, u8, v8, sdr_luma_stride, sdr_chroma_stride)
    return sdr, hdr


def natural_map(map_width, map_height, index=0):
    """A smooth u8 gain map spanning 0..255 (for apply-only tests)."""
    rng = _rng(index)
    f = _smooth_field(rng, map_height, map_width, 5)
    f = (f - f.min()) / max(float(f.max() - f.min()), 1e-6)
    return np.clip(np.rint(f * 255 + (rng.random((map_height, map_width)) - 0.5) * 6), 0, 255).astype(np.uint8)
