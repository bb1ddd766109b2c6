"""ctypes mirror of include/uhdr_b200.h (the C ABI of the hot path).

Enum and status values are numerically identical to the reference
(lib/include/ultrahdr/ultrahdr.h:36-120).
"""
import ctypes as C

import numpy as np

# ultrahdr_color_gamut (ultrahdr.h:36-42)
COLORGAMUT_UNSPECIFIED, COLORGAMUT_BT709, COLORGAMUT_P3, COLORGAMUT_BT2100 = -1, 0, 1, 2
# ultrahdr_transfer_function (ultrahdr.h:46-53)
TF_UNSPECIFIED, TF_LINEAR, TF_HLG, TF_PQ, TF_SRGB = -1, 0, 1, 2, 3
# ultrahdr_output_format (ultrahdr.h:56-64)
OUTPUT_UNSPECIFIED, OUTPUT_SDR, OUTPUT_HDR_LINEAR, OUTPUT_HDR_PQ, OUTPUT_HDR_HLG, OUTPUT_HDR_LINEAR_RGB_10BIT = -1, 0, 1, 2, 3, 4
# ultrahdr_pixel_format (ultrahdr.h:67-75)
PIX_FMT_UNSPECIFIED, PIX_FMT_P010, PIX_FMT_YUV420, PIX_FMT_MONOCHROME, PIX_FMT_RGBA8888, PIX_FMT_RGBAF16, PIX_FMT_RGBA1010102 = -1, 0, 1, 2, 3, 4, 5
# status_t (ultrahdr.h:91-120)
NO_ERROR = 0
UNKNOWN_ERROR = -1
ERROR_BAD_PTR = -10001
ERROR_UNSUPPORTED_WIDTH_HEIGHT = -10002
ERROR_INVALID_COLORGAMUT = -10003
ERROR_INVALID_STRIDE = -10004
ERROR_INVALID_TRANS_FUNC = -10005
ERROR_RESOLUTION_MISMATCH = -10006
ERROR_BAD_METADATA = -10010
ERROR_INVALID_QUALITY_FACTOR = -10007
ERROR_DECODE = -20002
ERROR_BUFFER_TOO_SMALL = -20004
ERROR_UNSUPPORTED_MAP_SCALE_FACTOR = -20008
ERROR_INSUFFICIENT_RESOURCE = -20009
ERROR_UNSUPPORTED_FEATURE = -30000
NOALLOC_MAX_IMAGES = 256  # UHDRB200_NOALLOC_MAX_IMAGES

APPLY_AUTO, APPLY_GENERIC, APPLY_TWO_TABLE, APPLY_Q_PIPE = 0, 1, 2, 3  # uhdrb200_test_set_apply_mode (include/uhdr_b200_test.h)
# uhdrb200_test_eval function selectors (include/uhdr_b200_test.h)
(FN_SRGB_INV_OETF, FN_HLG_OETF, FN_HLG_INV_OETF, FN_PQ_OETF, FN_PQ_INV_OETF, FN_FLOAT_TO_HALF, FN_TO_10BIT, FN_P_HLG, FN_P_PQ,
 FN_L_POW, FN_TOE_FIX, FN_ENCODE_GAIN) = range(12)
MODE_ANALYTIC, MODE_LUT = 0, 1  # uhdrb200_set_numeric_mode
SYNC_SPIN, SYNC_BLOCKING = 0, 1  # uhdrb200_set_sync_mode
MAP_SCALE = 4  # kMapDimensionScaleFactor, ultrahdr.h:213
SDR_WHITE_NITS, HLG_MAX_NITS, PQ_MAX_NITS = 203.0, 1000.0, 10000.0  # gainmapmath.h:34-36


class Image(C.Structure):
    """uhdrb200_image == ultrahdr_uncompressed_struct (ultrahdr.h:152-180)."""

    _fields_ = [
        ("data", C.c_void_p),
        ("width", C.c_size_t),
        ("height", C.c_size_t),
        ("colorGamut", C.c_int32),
        ("chroma_data", C.c_void_p),
        ("luma_stride", C.c_size_t),
        ("chroma_stride", C.c_size_t),
        ("pixelFormat", C.c_int32),
    ]


class Metadata(C.Structure):
    """uhdrb200_metadata == flattened ultrahdr_metadata_struct (ultrahdr.h:129-146)."""

    _fields_ = [
        ("version", C.c_char * 8),
        ("maxContentBoost", C.c_float),
        ("minContentBoost", C.c_float),
        ("gamma", C.c_float),
        ("offsetSdr", C.c_float),
        ("offsetHdr", C.c_float),
        ("hdrCapacityMin", C.c_float),
        ("hdrCapacityMax", C.c_float),
    ]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_}
        d["version"] = d["version"].decode()
        return d


class JpegComponent(C.Structure):
    """uhdrb200_jpeg_component (row N4): one colour component in the coefficient domain and its output plane."""

    _fields_ = [
        ("coeffs", C.c_void_p),
        ("quant", C.c_void_p),
        ("blocks_w", C.c_size_t),
        ("blocks_h", C.c_size_t),
        ("plane", C.c_void_p),
        ("width", C.c_size_t),
        ("height", C.c_size_t),
        ("stride", C.c_size_t),
    ]


def default_metadata(max_boost, min_boost=1.0):
    """What generateGainMap writes (ultrahdr.cpp:250-257)."""
    m = Metadata()
    m.version = b"1.0"
    m.maxContentBoost = max_boost
    m.minContentBoost = min_boost
    m.gamma = 1.0
    m.offsetSdr = 0.0
    m.offsetHdr = 0.0
    m.hdrCapacityMin = min_boost
    m.hdrCapacityMax = max_boost
    return m


def out_bytes_per_pixel(fmt):
    return {OUTPUT_HDR_LINEAR: 8, OUTPUT_HDR_PQ: 4, OUTPUT_HDR_HLG: 4, OUTPUT_HDR_LINEAR_RGB_10BIT: 6}.get(fmt, 0)


def _ptr(a, byte_offset=0):
    return C.c_void_p(a.ctypes.data + byte_offset)


def host_yuv420(buf, width, height, gamut, luma_stride=None, chroma_stride=None, chroma_offset=None):
    """Descriptor over a numpy uint8 buffer laid out Y | U | V (strides in pixels).

    Default layout is what the JPEG decoder hands applyGainMap (jpegr.cpp:790-799):
    luma stride = width, chroma right after luma, chroma stride = width/2, V plane at
    chroma + chroma_stride*height/2 (gainmapmath.cpp:568).
    """
    ls = luma_stride or width
    cs = chroma_stride or ls // 2
    off = ls * height if chroma_offset is None else chroma_offset
    im = Image()
    im.data = buf.ctypes.data
    im.width, im.height, im.colorGamut = width, height, gamut
    im.chroma_data = buf.ctypes.data + off
    im.luma_stride, im.chroma_stride = ls, cs
    im.pixelFormat = PIX_FMT_YUV420
    return im


def host_p010(buf, width, height, gamut, luma_stride=None, chroma_stride=None, chroma_offset=None):
    """Descriptor over a numpy uint16 buffer laid out Y | interleaved UV (strides in uint16 elements)."""
    ls = luma_stride or width
    cs = chroma_stride or ls
    off = ls * height if chroma_offset is None else chroma_offset  # in elements
    im = Image()
    im.data = buf.ctypes.data
    im.width, im.height, im.colorGamut = width, height, gamut
    im.chroma_data = buf.ctypes.data + 2 * off
    im.luma_stride, im.chroma_stride = ls, cs
    im.pixelFormat = PIX_FMT_P010
    return im


def host_map(buf, width, height):
    im = Image()
    im.data = buf.ctypes.data
    im.width, im.height, im.colorGamut = width, height, COLORGAMUT_UNSPECIFIED
    im.chroma_data = None
    im.luma_stride, im.chroma_stride = width, 0
    im.pixelFormat = PIX_FMT_MONOCHROME
    return im


def yuv420_size(width, height, luma_stride=None, chroma_stride=None):
    ls = luma_stride or width
    cs = chroma_stride or ls // 2
    return ls * height + 2 * cs * (height // 2)


def p010_size(width, height, luma_stride=None, chroma_stride=None):
    ls = luma_stride or width
    cs = chroma_stride or ls
    return ls * height + cs * (height // 2)


def alloc_output(fmt, width, height):
    return np.zeros(width * height * out_bytes_per_pixel(fmt), dtype=np.uint8)
