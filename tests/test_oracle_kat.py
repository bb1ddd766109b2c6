"""Known-answer tests of the reference's own unit tests (tests/gainmapmath_test.cpp), run against
the oracle (oracle/uhdr_oracle.c). Values, fixtures and epsilons are the reference's
(gainmapmath_test.cpp:29-31: 1e-4 comparison, 1e-2 luminance, 1/510 YUV). When oracle/_ref is
present the same KATs are also run against the unmodified reference build, which proves the
harness itself."""
import ctypes as C

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi

EPS, LUM_EPS, YUV_EPS = 1e-4, 1e-2, 1.0 / 510.0
BLACK, WHITE = (0.0, 0.0, 0.0), (1.0, 1.0, 1.0)
RED, GREEN, BLUE = (1.0, 0.0, 0.0), (0.0, 1.0, 0.0), (0.0, 0.0, 1.0)
YUV_BLACK, YUV_WHITE = (0.0, 0.0, 0.0), (1.0, 0.0, 0.0)
# gainmapmath_test.cpp:62-72
YUV_PRIMARIES = {
    0: ((0.2126, -0.11457, 0.5), (0.7152, -0.38543, -0.45415), (0.0722, 0.5, -0.04585)),
    1: ((0.299, -0.16874, 0.5), (0.587, -0.33126, -0.41869), (0.114, 0.5, -0.08131)),
    2: ((0.2627, -0.13963, 0.5), (0.6780, -0.36037, -0.45979), (0.0593, 0.5, -0.04021)),
}
LUMA = {0: (0.2126, 0.7152, 0.0722), 1: (0.20949, 0.72160, 0.06891), 2: (0.2627, 0.6780, 0.0593)}


@pytest.fixture(params=["oracle", "reference"])
def impl(request):
    if request.param == "oracle":
        return request.getfixturevalue("orc")
    return request.getfixturevalue("ref")


def f32(x):
    return float(np.float32(x))


# ---- 4x4 fixtures, gainmapmath_test.cpp:106-267
def yuv420_fixture():
    px = np.array([0x00, 0x10, 0x20, 0x30, 0x01, 0x11, 0x21, 0x31, 0x02, 0x12, 0x22, 0x32, 0x03, 0x13, 0x23, 0x33,
                   0xA0, 0xA1, 0xA2, 0xA3, 0xB0, 0xB1, 0xB2, 0xB3], dtype=np.uint8)
    return px, abi.host_yuv420(px, 4, 4, 0, 4, 2)


def p010_fixture():
    vals = [0x00, 0x10, 0x20, 0x30, 0x01, 0x11, 0x21, 0x31, 0x02, 0x12, 0x22, 0x32, 0x03, 0x13, 0x23, 0x33,
            0xA0, 0xB0, 0xA1, 0xB1, 0xA2, 0xB2, 0xA3, 0xB3]
    px = (np.array(vals, dtype=np.uint16) << 6).astype(np.uint16)
    return px, abi.host_p010(px, 4, 4, 0, 4, 4)


def map_fixture():
    px = np.array([0x00, 0x10, 0x20, 0x30, 0x01, 0x11, 0x21, 0x31, 0x02, 0x12, 0x22, 0x32, 0x03, 0x13, 0x23, 0x33], dtype=np.uint8)
    return px, abi.host_map(px, 4, 4)


def yuv420_color(y, u, v):  # gainmapmath_test.cpp:33-36
    k = np.float32(1 / 255.0)
    return np.array([np.float32(y) * k, np.float32(u - 128) * k, np.float32(v - 128) * k], dtype=np.float32)


def p010_color(y, u, v):  # gainmapmath_test.cpp:38-42
    return np.array([np.float32(y - 64) * np.float32(1 / 876.0), np.float32(u - 64) * np.float32(1 / 896.0) - np.float32(0.5),
                     np.float32(v - 64) * np.float32(1 / 896.0) - np.float32(0.5)], dtype=np.float32)


def expected_colors(fn):
    ys = [[0x00, 0x10, 0x20, 0x30], [0x01, 0x11, 0x21, 0x31], [0x02, 0x12, 0x22, 0x32], [0x03, 0x13, 0x23, 0x33]]
    us, vs = [[0xA0, 0xA1], [0xA2, 0xA3]], [[0xB0, 0xB1], [0xB2, 0xB3]]
    return [[fn(ys[y][x], us[y // 2][x // 2], vs[y // 2][x // 2]) for x in range(4)] for y in range(4)]


# ---- transfer functions
def test_srgb_transfer_function(impl):  # :467-473
    assert impl.c_srgbInvOetf(0.0) == 0.0
    assert abs(impl.c_srgbInvOetf(0.02) - 0.00154) < EPS
    assert abs(impl.c_srgbInvOetf(0.04045) - 0.00313) < EPS
    assert abs(impl.c_srgbInvOetf(0.5) - 0.21404) < EPS
    assert impl.c_srgbInvOetf(1.0) == pytest.approx(1.0, rel=4e-7)


def test_hlg(impl):  # :752-782
    assert impl.c_hlgOetf(0.0) == 0.0
    for x, y in ((0.04167, 0.35357), (0.08333, 0.5), (0.5, 0.87164)):
        assert abs(impl.c_hlgOetf(x) - y) < EPS
        assert abs(impl.c_hlgInvOetf(impl.c_hlgOetf(x)) - x) < EPS
    assert impl.c_hlgOetf(1.0) == pytest.approx(1.0, rel=4e-7)
    assert impl.c_hlgInvOetf(0.0) == 0.0
    for x, y in ((0.25, 0.02083), (0.5, 0.08333), (0.75, 0.26496)):
        assert abs(impl.c_hlgInvOetf(x) - y) < EPS
    assert impl.c_hlgInvOetf(1.0) == pytest.approx(1.0, rel=4e-7)


def test_pq(impl):  # :784-806, 941-947
    assert impl.c_pqOetf(0.0) == 0.0
    for x, y in ((0.01, 0.50808), (0.5, 0.92655), (0.99, 0.99895)):
        assert abs(impl.c_pqOetf(x) - y) < EPS
        assert abs(impl.c_pqInvOetf(impl.c_pqOetf(x)) - x) < EPS
    assert impl.c_pqOetf(1.0) == pytest.approx(1.0, rel=4e-7)
    assert impl.c_pqInvOetf(0.0) == 0.0
    for x, y in ((0.01, 2.31017e-7), (0.5, 0.00922), (0.99, 0.90903)):
        assert abs(impl.c_pqInvOetf(x) - y) < EPS
    assert impl.c_pqInvOetf(1.0) == pytest.approx(1.0, rel=4e-7)


@pytest.mark.parametrize("fn,n", [("srgbInvOetf", 1024), ("hlgInvOetf", 4096), ("pqInvOetf", 4096), ("hlgOetf", 65536), ("pqOetf", 65536)])
def test_lut_equals_analytic_at_nodes(impl, fn, n):  # :808-841, table sizes gainmapmath.h:268-369
    step = max(1, n // 512)
    for idx in list(range(0, n, step)) + [n - 1]:
        v = f32(np.float32(idx) / np.float32(n - 1))
        assert getattr(impl, "c_" + fn)(v) == getattr(impl, "c_" + fn + "LUT")(v)


# ---- luminance, YUV <-> RGB (gainmapmath_test.cpp:416-466, 475-590)
@pytest.mark.parametrize("g", [0, 1, 2])
def test_luminance_and_yuv_matrices(impl, g):
    for rgb, want in ((BLACK, 0.0), (WHITE, 1.0), (RED, LUMA[g][0]), (GREEN, LUMA[g][1]), (BLUE, LUMA[g][2])):
        assert impl.c_luminance(g, *rgb) == pytest.approx(want, rel=4e-7, abs=1e-9)
    to_rgb, to_yuv = impl.vec3("yuvToRgb", g), impl.vec3("rgbToYuv", g)
    for yuv, rgb in zip((YUV_BLACK, YUV_WHITE) + YUV_PRIMARIES[g], (BLACK, WHITE, RED, GREEN, BLUE)):
        assert np.abs(to_rgb(yuv) - np.array(rgb)).max() < EPS
        assert np.abs(to_yuv(rgb) - np.array(yuv)).max() < EPS
        assert np.abs(to_rgb(to_yuv(rgb)) - np.array(rgb)).max() < EPS


def test_yuv_to_yuv_conversions(impl):  # :592-692: a round trip through another encoding is the identity within 1/510
    back = {0: 2, 1: 4, 2: 0, 3: 5, 4: 1, 5: 3}  # idx -> its inverse (709->601->709, 709->2100->709, ...)
    rng = np.random.Generator(np.random.PCG64(1))
    for _ in range(50):
        v = np.array([rng.uniform(0, 1), rng.uniform(-0.5, 0.5), rng.uniform(-0.5, 0.5)], dtype=np.float32)
        for fwd, inv in back.items():
            out = impl.vec3("yuvToYuv", inv)(impl.vec3("yuvToYuv", fwd)(v))
            assert np.abs(out - v).max() < YUV_EPS


def test_transform_yuv420(impl):  # :694-750: transformYuv420 == per-pixel conversion + 2x2 chroma average
    px, img = yuv420_fixture()
    for s, d, idx in ((0, 1, 0), (0, 2, 1), (1, 0, 2), (1, 2, 3), (2, 0, 4), (2, 1, 5)):
        buf = px.copy()
        im = abi.host_yuv420(buf, 4, 4, s, 4, 2)
        assert impl.convert_yuv(im, s, d) == 0
        conv = impl.vec3("yuvToYuv", idx)
        cols = expected_colors(yuv420_color)
        for cy in range(2):
            for cx in range(2):
                outs = [conv(cols[cy * 2 + dy][cx * 2 + dx]) for dy in range(2) for dx in range(2)]
                for k, (dy, dx) in enumerate(((0, 0), (0, 1), (1, 0), (1, 1))):
                    want_y = int(np.clip(np.float32(outs[k][0]) * np.float32(255.0) + np.float32(0.5), 0, 255))
                    assert abs(int(buf[(cy * 2 + dy) * 4 + cx * 2 + dx]) - want_y) <= 1
                mu = np.mean([o[1] for o in outs]) * 255.0 + 128.0
                mv = np.mean([o[2] for o in outs]) * 255.0 + 128.0
                assert abs(int(buf[16 + cy * 2 + cx]) - mu) <= 1.0
                assert abs(int(buf[20 + cy * 2 + cx]) - mv) <= 1.0


# ---- encodeGain / applyGain (gainmapmath_test.cpp:980-1113)
def test_encode_gain_exact_bytes(impl):
    cases = [
        ((0.25, 4.0), [((0, 0), 127), ((0, 1), 127), ((1, 0), 0), ((0.5, 0), 0), ((1, 1), 127), ((1, 4), 255), ((1, 5), 255),
                       ((4, 1), 0), ((4, 0.5), 0), ((1, 2), 191), ((2, 1), 63)]),
        ((0.5, 2.0), [((1, 2), 255), ((2, 1), 0), ((1, 1.41421), 191), ((1.41421, 1), 63)]),
        ((0.125, 8.0), [((1, 8), 255), ((8, 1), 0), ((1, 2.82843), 191), ((2.82843, 1), 63)]),
        ((1.0, 8.0), [((0, 0), 0), ((1, 0), 0), ((1, 1), 0), ((1, 8), 255), ((1, 4), 170), ((1, 2), 85)]),
        ((0.5, 8.0), [((0, 0), 63), ((1, 0), 0), ((1, 1), 63), ((1, 8), 255), ((1, 4), 191), ((1, 2), 127), ((1, 0.7071), 31), ((1, 0.5), 0)]),
    ]
    for (mn, mx), rows in cases:
        for (ys, yh), want in rows:
            assert impl.c_encodeGain(f32(ys), f32(yh), f32(mn), f32(mx)) == want, ((mn, mx), (ys, yh))


def test_apply_gain(impl):
    w = np.array(WHITE, dtype=np.float32)
    table = [((0.25, 4.0), [(0.0, 1 / 4), (0.25, 1 / 2), (0.5, 1.0), (0.75, 2.0), (1.0, 4.0)]),
             ((0.5, 2.0), [(0.0, 1 / 2), (0.25, 1 / 1.41421), (0.5, 1.0), (0.75, 1.41421), (1.0, 2.0)]),
             ((0.125, 8.0), [(0.0, 1 / 8), (0.25, 1 / 2.82843), (0.5, 1.0), (0.75, 2.82843), (1.0, 8.0)]),
             ((1.0, 8.0), [(0.0, 1.0), (1 / 3, 2.0), (2 / 3, 4.0), (1.0, 8.0)]),
             ((0.5, 8.0), [(0.0, 0.5), (0.25, 1.0), (0.5, 2.0), (0.75, 4.0), (1.0, 8.0)])]
    for (mn, mx), rows in table:
        for g, scale in rows:
            assert np.abs(impl.apply_gain(w, f32(g), mn, mx) - w * scale).max() < EPS * max(1.0, scale)
            assert np.abs(impl.apply_gain(np.zeros(3, np.float32), f32(g), mn, mx)).max() == 0.0
    e = np.array([0.0, 0.5, 1.0], dtype=np.float32)
    for g, scale in ((0.0, 0.25), (0.25, 0.5), (0.5, 1.0), (0.75, 2.0), (1.0, 4.0)):
        assert np.abs(impl.apply_gain(e, g, 0.25, 4.0) - e * scale).max() < EPS * 4
    for rgb in (BLACK, WHITE, RED, GREEN, BLUE, tuple(e)):  # displayBoost == maxContentBoost is the identity
        a = impl.apply_gain(np.array(rgb, np.float32), 1.0, 0.25, 4.0)
        b = impl.apply_gain(np.array(rgb, np.float32), 1.0, 0.25, 4.0, 4.0)
        assert np.array_equal(a, b)


def test_apply_gain_lut(impl):  # :843-939 (condensed): LUT == analytic within 1e-4 relative steps of the 1024-entry table
    for mn, mx in ((0.25, 4.0), (0.5, 2.0), (0.125, 8.0), (1.0, 8.0), (0.5, 8.0)):
        for idx in range(0, 1024, 17):
            g = f32(np.float32(idx) / np.float32(1023.0))
            for rgb in (WHITE, RED, (0.0, 0.5, 1.0)):
                a = impl.apply_gain_lut(np.array(rgb, np.float32), g, mn, mx, mx)
                b = impl.apply_gain(np.array(rgb, np.float32), g, mn, mx)
                assert np.abs(a - b).max() <= 1e-4 * mx


# ---- pixel accessors and samplers (gainmapmath_test.cpp:1115-1227)
def test_get_pixels(impl):
    _, y420 = yuv420_fixture()
    _, p010 = p010_fixture()
    c420, c010 = expected_colors(yuv420_color), expected_colors(p010_color)
    for y in range(4):
        for x in range(4):
            assert np.abs(impl.pixel("getYuv420Pixel", y420, x, y) - c420[y][x]).max() < EPS
            assert np.abs(impl.pixel("getP010Pixel", p010, x, y) - c010[y][x]).max() < EPS


def test_sample_between_min_max(impl):
    for fixture, fn, colors in ((yuv420_fixture, "sampleYuv420", expected_colors(yuv420_color)),
                                (p010_fixture, "sampleP010", expected_colors(p010_color))):
        _, img = fixture()
        for y in range(2):
            for x in range(2):
                blk = np.array([colors[y * 2 + dy][x * 2 + dx] for dy in range(2) for dx in range(2)])
                got = impl.pixel(fn, img, 2, x, y)
                assert (got >= blk.min(axis=0) - 1e-6).all() and (got <= blk.max(axis=0) + 1e-6).all()


def test_sample_map_idw_equals_float_path(impl):
    px, img = map_fixture()
    vals = (px.astype(np.float32) / np.float32(255.0)).reshape(4, 4)
    for y in range(8):
        for x in range(8):
            xb, yb = x // 2, y // 2
            nb = [vals[yy][xx] for yy in (yb, min(yb + 1, 3)) for xx in (xb, min(xb + 1, 3))]
            a = impl.c_sampleMapIdw(C.byref(img), 2, x, y)
            b = impl.c_sampleMapFloat(C.byref(img), 2.0, x, y)
            assert min(nb) - 1e-7 <= b <= max(nb) + 1e-7
            assert np.float32(a) == np.float32(b)


# ---- packers (gainmapmath_test.cpp:1229-1262)
def test_color_to_rgba1010102(impl):
    assert impl.c_colorToRgba1010102(*BLACK) == 0x3 << 30
    assert impl.c_colorToRgba1010102(*WHITE) == 0xFFFFFFFF
    assert impl.c_colorToRgba1010102(*RED) == (0x3 << 30 | 0x3FF)
    assert impl.c_colorToRgba1010102(*GREEN) == (0x3 << 30 | 0x3FF << 10)
    assert impl.c_colorToRgba1010102(*BLUE) == (0x3 << 30 | 0x3FF << 20)
    q = [int(np.float32(v) * np.float32(1023.0)) for v in (0.1, 0.2, 0.3)]
    assert impl.c_colorToRgba1010102(f32(0.1), f32(0.2), f32(0.3)) == (0x3 << 30 | q[0] | q[1] << 10 | q[2] << 20)


def test_color_to_rgba_f16_and_float_to_half(impl):
    assert impl.c_colorToRgbaF16(*BLACK) == 0x3C00 << 48
    assert impl.c_colorToRgbaF16(*WHITE) == 0x3C003C003C003C00
    assert impl.c_colorToRgbaF16(*RED) == (0x3C00 << 48 | 0x3C00)
    assert impl.c_colorToRgbaF16(*GREEN) == (0x3C00 << 48 | 0x3C00 << 16)
    assert impl.c_colorToRgbaF16(*BLUE) == (0x3C00 << 48 | 0x3C00 << 32)
    assert impl.c_colorToRgbaF16(f32(0.1), f32(0.2), f32(0.3)) == 0x3C0034CD32662E66
    for v, want in ((0.1, 0x2E66), (0.0, 0x0), (1.0, 0x3C00), (-1.0, 0xBC00), (3.4028234663852886e38, 0x7FFF),
                    (-3.4028234663852886e38, 0xFFFF), (1.1754943508222875e-38, 0x0)):
        assert impl.c_floatToHalf(f32(v)) == want


# ---- luminance chains and ApplyMap (gainmapmath_test.cpp:1264-1386)
def _chain(impl, yuv, yuv_gamut, inv, luma_gamut, scale):
    rgb = impl.vec3("yuvToRgb", yuv_gamut)(yuv)
    lin = [getattr(impl, "c_" + inv)(float(v)) if inv else float(v) for v in rgb]
    return impl.c_luminance(luma_gamut, *lin) * scale


def test_generate_map_luminance_chains(impl):
    for lg in (0, 1, 2):
        assert _chain(impl, YUV_BLACK, 0, "srgbInvOetf", lg, 203.0) == 0.0
        assert _chain(impl, YUV_WHITE, 0, "srgbInvOetf", lg, 203.0) == pytest.approx(203.0, rel=1e-6)
        for yuv, rgb in zip(YUV_PRIMARIES[0], (RED, GREEN, BLUE)):
            assert abs(_chain(impl, yuv, 0, "srgbInvOetf", lg, 203.0) - impl.c_luminance(lg, *rgb) * 203.0) < LUM_EPS
    for inv, nits in (("hlgInvOetf", 1000.0), ("pqInvOetf", 10000.0)):
        assert _chain(impl, YUV_BLACK, 2, inv, 2, nits) == 0.0
        assert _chain(impl, YUV_WHITE, 2, inv, 2, nits) == pytest.approx(nits, rel=1e-6)
        for yuv, rgb in zip(YUV_PRIMARIES[2], (RED, GREEN, BLUE)):
            # the reference's LuminanceEpsilon (1e-2) is applied at its scale; PQ's 10000 nits amplifies the
            # 5-decimal YUV literals, hence the relative form
            assert abs(_chain(impl, yuv, 2, inv, 2, nits) - impl.c_luminance(2, *rgb) * nits) < max(LUM_EPS, 2e-3 * nits)


def test_apply_map_recover(impl):
    def recover(yuv, gain, mn, mx):
        rgb = impl.vec3("yuvToRgb", 0)(yuv)
        lin = np.array([impl.c_srgbInvOetf(float(v)) for v in rgb], dtype=np.float32)
        return impl.apply_gain(lin, f32(gain), mn, mx)

    s8 = float(np.sqrt(np.float32(8.0)))
    for gain, scale in ((1.0, 8.0), (0.75, s8), (0.5, 1.0), (0.25, 1 / s8), (0.0, 1 / 8.0)):
        assert np.abs(recover(YUV_WHITE, gain, 0.125, 8.0) - np.array(WHITE) * scale).max() < 1e-5 * max(scale, 1)
        assert np.abs(recover(YUV_BLACK, gain, 0.125, 8.0)).max() == 0.0
        for yuv, rgb in zip(YUV_PRIMARIES[0], (RED, GREEN, BLUE)):
            assert np.abs(recover(yuv, gain, 0.125, 8.0) - np.array(rgb) * scale).max() < EPS * 50 * max(scale, 1)
    for gain, scale in ((1.0, 8.0), (2 / 3, 4.0), (1 / 3, 2.0), (0.0, 1.0)):
        assert np.abs(recover(YUV_WHITE, gain, 1.0, 8.0) - scale).max() < 1e-5 * scale
    for gain, scale in ((1.0, 8.0), (0.75, 4.0), (0.5, 2.0), (0.25, 1.0), (0.0, 0.5)):
        assert np.abs(recover(YUV_WHITE, gain, 0.5, 8.0) - scale).max() < 1e-5 * scale
