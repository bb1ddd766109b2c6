"""ctypes bindings of the CPU checkers — TEST INFRASTRUCTURE ONLY.

Two interchangeable implementations of the same C surface:
  * ``CpuImpl.oracle()``     -> oracle/liboracle.so      (our plain-C restatement, prefix ``orc_``)
  * ``CpuImpl.reference()``  -> oracle/_ref/libuhdr_ref.so (the UNMODIFIED reference hot path
                                 compiled by oracle/Makefile, prefix ``ref_``); ``lut=True`` picks
                                 the -DUSE_*_LUT=1 variant.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import
this module; the product package never does.
"""
import ctypes as C
import os

import numpy as np

from libultrahdr_dev_b200 import abi

HERE = os.path.dirname(os.path.abspath(__file__))
P_IMG = C.POINTER(abi.Image)
P_MD = C.POINTER(abi.Metadata)
P_F = C.POINTER(C.c_float)


def _f3(a):
    return (C.c_float * 3)(*[float(v) for v in a])


class CpuImpl:
    def __init__(self, path, prefix, name):
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self.name = name
        self.path = path
        L = self._fn
        L("generate_gainmap", C.c_int, [P_IMG, P_IMG, C.c_int, P_MD, P_IMG, C.c_int, C.c_void_p, C.c_size_t])
        L("apply_gainmap", C.c_int, [P_IMG, P_IMG, P_MD, C.c_int, C.c_float, P_IMG])
        L("tonemap", C.c_int, [P_IMG, P_IMG])
        L("convert_yuv", C.c_int, [P_IMG, C.c_int, C.c_int])
        L("is_lut_variant", C.c_int, [])
        for n in ("srgbInvOetf", "hlgOetf", "hlgInvOetf", "pqOetf", "pqInvOetf"):
            L(n, C.c_float, [C.c_float])
            L(n + "LUT", C.c_float, [C.c_float])
        L("luminance", C.c_float, [C.c_int, C.c_float, C.c_float, C.c_float])
        L("yuvToRgb", None, [C.c_int, P_F, P_F])
        L("rgbToYuv", None, [C.c_int, P_F, P_F])
        L("hdrConversion", C.c_int, [C.c_int, C.c_int, P_F, P_F])
        L("yuvToYuv", None, [C.c_int, P_F, P_F])
        L("encodeGain", C.c_int, [C.c_float, C.c_float, C.c_float, C.c_float])
        L("applyGain", None, [P_F, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, P_F])
        L("applyGainLUT", None, [P_F, C.c_float, C.c_float, C.c_float, C.c_float, P_F])
        for n in ("getYuv420Pixel", "getP010Pixel"):
            L(n, None, [P_IMG, C.c_size_t, C.c_size_t, P_F])
        for n in ("sampleYuv420", "sampleP010"):
            L(n, None, [P_IMG, C.c_size_t, C.c_size_t, C.c_size_t, P_F])
        L("sampleMapIdw", C.c_float, [P_IMG, C.c_size_t, C.c_size_t, C.c_size_t])
        L("sampleMapFloat", C.c_float, [P_IMG, C.c_float, C.c_size_t, C.c_size_t])
        L("idwTable", None, [C.c_int, C.c_int, P_F])
        L("colorToRgba1010102", C.c_uint32, [C.c_float, C.c_float, C.c_float])
        L("colorToRgbaF16", C.c_uint64, [C.c_float, C.c_float, C.c_float])
        L("floatToHalf", C.c_uint16, [C.c_float])
        # row N3 exists in the restatement only (the reference's routine lives in its example app: ref_app below)
        if hasattr(self.lib, self.prefix + "yuv420_to_rgba8888"):
            L("yuv420_to_rgba8888", C.c_int, [P_IMG, P_IMG])
            L("copy_yuv420", C.c_int, [P_IMG, P_IMG])

    def _fn(self, name, restype, argtypes):
        f = getattr(self.lib, self.prefix + name)
        f.restype = restype
        f.argtypes = argtypes
        setattr(self, "c_" + name, f)

    # ---- constructors -------------------------------------------------------------------
    @staticmethod
    def oracle(lut=False):
        p = os.path.join(HERE, "liboracle.so")
        if not os.path.exists(p):
            raise FileNotFoundError(p + " missing: run `make -C oracle oracle` (or __graft_entry__.build())")
        impl = CpuImpl(p, "orc_", "oracle-lut" if lut else "oracle")
        impl.lib.orc_set_lut_mode(1 if lut else 0)
        return impl

    @staticmethod
    def reference_path(lut=False):
        return os.path.join(HERE, "_ref", "libuhdr_ref_lut.so" if lut else "libuhdr_ref.so")

    @staticmethod
    def reference(lut=False):
        p = CpuImpl.reference_path(lut)
        if not os.path.exists(p):
            raise FileNotFoundError(p + " missing: run `make -C oracle ref` where /root/reference exists")
        return CpuImpl(p, "ref_", "reference-lut" if lut else "reference")

    @staticmethod
    def have_reference(lut=False):
        return os.path.exists(CpuImpl.reference_path(lut))

    # ---- the three path entry points (numpy in / numpy out) --------------------------------
    def generate_gainmap(self, sdr_img, hdr_img, hdr_tf, sdr_is_601=False):
        """-> (status, map ndarray[h/4, w/4] u8 or None, Metadata, dest Image)"""
        mw, mh = sdr_img.width // abi.MAP_SCALE, sdr_img.height // abi.MAP_SCALE
        buf = np.zeros(max(mw * mh, 1), dtype=np.uint8)
        md, dest = abi.Metadata(), abi.Image()
        st = self.c_generate_gainmap(C.byref(sdr_img), C.byref(hdr_img), hdr_tf, C.byref(md), C.byref(dest),
                                     1 if sdr_is_601 else 0, buf.ctypes.data, buf.size)
        if st != 0:
            return st, None, md, dest
        return st, buf[: mw * mh].reshape(mh, mw), md, dest

    def apply_gainmap(self, sdr_img, map_img, md, fmt, max_display_boost):
        """-> (status, flat u8 output buffer, dest Image)"""
        out = abi.alloc_output(fmt, sdr_img.width, sdr_img.height)
        if out.size == 0:
            out = np.zeros(sdr_img.width * sdr_img.height * 8, dtype=np.uint8)
        dest = abi.Image()
        dest.data = out.ctypes.data
        st = self.c_apply_gainmap(C.byref(sdr_img), C.byref(map_img), C.byref(md), fmt, max_display_boost, C.byref(dest))
        return st, out, dest

    def tonemap(self, src_img, dest_img):
        return self.c_tonemap(C.byref(src_img), C.byref(dest_img))

    def convert_yuv(self, img, src_enc, dst_enc):
        return self.c_convert_yuv(C.byref(img), src_enc, dst_enc)

    def yuv420_to_rgba8888(self, sdr_img, out_stride=0):
        """-> (status, uint32 array of height x stride)."""
        import numpy as np

        stride = max(int(out_stride), int(sdr_img.width))
        out = np.zeros(max(int(sdr_img.height) * stride, 1), dtype=np.uint32)
        dest = abi.Image()
        dest.data = out.ctypes.data
        dest.luma_stride = out_stride
        st = self.c_yuv420_to_rgba8888(C.byref(sdr_img), C.byref(dest))
        return st, out[: int(sdr_img.height) * stride].reshape(int(sdr_img.height), stride), dest

    def copy_yuv420(self, src_img, dst_img):
        return self.c_copy_yuv420(C.byref(src_img), C.byref(dst_img))

    # ---- L1 conveniences ------------------------------------------------------------------
    def vec3(self, fn, *lead):
        def call(v):
            o = (C.c_float * 3)()
            getattr(self, "c_" + fn)(*lead, _f3(v), o)
            return np.array(list(o), dtype=np.float32)
        return call

    def apply_gain(self, rgb, gain, mn, mx, db=None):
        o = (C.c_float * 3)()
        self.c_applyGain(_f3(rgb), gain, mn, mx, db if db is not None else 0.0, 0 if db is None else 1, o)
        return np.array(list(o), dtype=np.float32)

    def apply_gain_lut(self, rgb, gain, mn, mx, db):
        o = (C.c_float * 3)()
        self.c_applyGainLUT(_f3(rgb), gain, mn, mx, db, o)
        return np.array(list(o), dtype=np.float32)

    def pixel(self, fn, img, *args):
        o = (C.c_float * 3)()
        getattr(self, "c_" + fn)(C.byref(img), *args, o)
        return np.array(list(o), dtype=np.float32)

    def idw_table(self, scale, table):
        o = (C.c_float * (scale * scale * 4))()
        self.c_idwTable(scale, table, o)
        return np.array(list(o), dtype=np.float32)


REF_APP_PATH = os.path.join(HERE, "_ref", "libuhdr_ref_app.so")


def have_ref_app():
    return os.path.exists(REF_APP_PATH)


def ref_app_yuv420_to_rgba8888(y, u, v, w, h, y_stride, u_stride, v_stride):
    """The reference's own UltraHdrAppInput::convertYuv420ToRGBImage (examples/ultrahdr_app.cpp:682-734),
    compiled unmodified into oracle/_ref/libuhdr_ref_app.so. -> uint32 array (h, w)."""
    import numpy as np

    lib = C.CDLL(REF_APP_PATH)
    fn = lib.ref_app_yuv420_to_rgba8888
    fn.restype = C.c_int
    fn.argtypes = [C.c_void_p] * 3 + [C.c_int] * 5 + [C.c_void_p]
    out = np.zeros(w * h, dtype=np.uint32)
    rc = fn(y.ctypes.data, u.ctypes.data, v.ctypes.data, w, h, y_stride, u_stride, v_stride, out.ctypes.data)
    if rc != 0:
        raise RuntimeError("reference app routine failed")
    return out.reshape(h, w)

