"""Python host-side mirror of the reference's hot-path interface, calling the C ABI.

`UltraHdr.generateGainMap / applyGainMap / toneMap` keep the reference's names, argument order and
status codes (lib/include/ultrahdr/ultrahdr.h:348-381); `JpegR.convertYuv` mirrors
lib/src/jpegr.cpp:1132. Python is used only by tests/ and bench.py — products bind the C ABI
(include/uhdr_b200.h) directly, see INTEGRATION.md.

No CPU fallback exists: a missing library raises at load time and a CUDA failure raises
UltraHdrCudaError (the C ABI returns ERROR_ULTRAHDR_INSUFFICIENT_RESOURCE with the reason in
uhdrb200_last_error()).
"""
import ctypes as C
import os

from . import abi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libuhdr_b200.so")

P_IMG = C.POINTER(abi.Image)
P_MD = C.POINTER(abi.Metadata)

# every symbol include/uhdr_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "uhdrb200_abi_version": (C.c_int, []),
    "uhdrb200_device_count": (C.c_int, []),
    "uhdrb200_init": (C.c_int, [C.c_int]),
    "uhdrb200_shutdown": (None, []),
    "uhdrb200_last_error": (C.c_char_p, []),
    "uhdrb200_host_alloc": (C.c_void_p, [C.c_size_t]),
    "uhdrb200_host_free": (None, [C.c_void_p]),
    "uhdrb200_free": (None, [C.c_void_p]),
    "uhdrb200_launch_count": (C.c_uint64, []),
    "uhdrb200_set_sync_mode": (C.c_int, [C.c_int]),
    "uhdrb200_set_numeric_mode": (C.c_int, [C.c_int]),
    "uhdrb200_get_numeric_mode": (C.c_int, []),
    "uhdrb200_generate_gainmap_host": (C.c_int, [P_IMG, P_IMG, C.c_int, P_MD, P_IMG, C.c_int]),
    "uhdrb200_generate_gainmap_host_into": (C.c_int, [P_IMG, P_IMG, C.c_int, P_MD, P_IMG, C.c_int, C.c_void_p, C.c_size_t]),
    "uhdrb200_generate_gainmap_dev": (C.c_int, [P_IMG, P_IMG, C.c_int, P_MD, P_IMG, C.c_int, C.c_void_p]),
    "uhdrb200_generate_gainmap_batch_dev": (C.c_int, [C.c_int, P_IMG, P_IMG, C.c_int, P_MD, P_IMG, C.c_int, C.c_void_p]),
    "uhdrb200_generate_gainmap_batch_dev_noalloc": (C.c_int, [C.c_int, P_IMG, P_IMG, C.c_int, P_MD, P_IMG, C.c_int, C.c_void_p]),
    "uhdrb200_apply_gainmap_host": (C.c_int, [P_IMG, P_IMG, P_MD, C.c_int, C.c_float, P_IMG]),
    "uhdrb200_apply_gainmap_dev": (C.c_int, [P_IMG, P_IMG, P_MD, C.c_int, C.c_float, P_IMG, C.c_void_p]),
    "uhdrb200_apply_gainmap_batch_dev": (C.c_int, [C.c_int, P_IMG, P_IMG, P_MD, C.c_int, C.c_float, P_IMG, C.c_void_p]),
    "uhdrb200_tonemap_host": (C.c_int, [P_IMG, P_IMG]),
    "uhdrb200_tonemap_dev": (C.c_int, [P_IMG, P_IMG, C.c_void_p]),
    "uhdrb200_tonemap_batch_dev": (C.c_int, [C.c_int, P_IMG, P_IMG, C.c_void_p]),
    "uhdrb200_tonemap_generate_gainmap_host_into": (C.c_int, [P_IMG, C.c_int, C.c_int, P_IMG, P_MD, P_IMG, C.c_void_p, C.c_size_t]),
    "uhdrb200_tonemap_generate_gainmap_dev": (C.c_int, [P_IMG, C.c_int, C.c_int, P_IMG, P_MD, P_IMG, C.c_void_p]),
    "uhdrb200_tonemap_generate_gainmap_batch_dev": (C.c_int, [C.c_int, P_IMG, C.c_int, C.c_int, P_IMG, P_MD, P_IMG, C.c_void_p]),
    "uhdrb200_convert_yuv_host": (C.c_int, [P_IMG, C.c_int, C.c_int]),
    "uhdrb200_convert_yuv_dev": (C.c_int, [P_IMG, C.c_int, C.c_int, C.c_void_p]),
    "uhdrb200_yuv420_to_rgba8888_host": (C.c_int, [P_IMG, P_IMG]),
    "uhdrb200_yuv420_to_rgba8888_dev": (C.c_int, [P_IMG, P_IMG, C.c_void_p]),
    "uhdrb200_yuv420_to_rgba8888_batch_dev": (C.c_int, [C.c_int, P_IMG, P_IMG, C.c_void_p]),
    "uhdrb200_jpeg_idct_host": (C.c_int, [C.c_int, C.POINTER(abi.JpegComponent)]),
    "uhdrb200_jpeg_idct_dev": (C.c_int, [C.c_int, C.POINTER(abi.JpegComponent), C.c_void_p]),
    "uhdrb200_jpeg_get_info": (C.c_int, [C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_size_t), C.POINTER(C.c_int)]),
    "uhdrb200_jpeg_decode_host": (C.c_int, [C.c_void_p, C.c_size_t, P_IMG]),
    "uhdrb200_jpeg_decode_dev": (C.c_int, [C.c_void_p, C.c_size_t, P_IMG, C.c_void_p]),
    "uhdrb200_jpeg_encode_host": (C.c_int, [P_IMG, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "uhdrb200_jpeg_encode_dev": (C.c_int, [P_IMG, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.c_void_p]),
    "uhdrb200_copy_yuv420_host": (C.c_int, [P_IMG, P_IMG]),
    "uhdrb200_copy_yuv420_dev": (C.c_int, [P_IMG, P_IMG, C.c_void_p]),
}


# test hooks (include/uhdr_b200_test.h): not part of the drop-in boundary
TEST_SYMBOLS = {
    "uhdrb200_test_set_apply_mode": (C.c_int, [C.c_int]),
    "uhdrb200_test_eval": (C.c_int, [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "uhdrb200_test_jpeg_headers": (C.c_size_t, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "uhdrb200_test_jpeg_stuff": (C.c_size_t, [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]),
    "uhdrb200_test_jpeg_destuff": (C.c_size_t, [C.c_void_p, C.c_size_t, C.c_void_p, C.POINTER(C.c_int), C.c_uint, C.POINTER(C.c_uint), C.POINTER(C.c_uint)]),
    "uhdrb200_test_jpeg_subseq_bytes": (C.c_size_t, []),
}


class UltraHdrCudaError(RuntimeError):
    pass


_lib = None


def load_library(path=None):
    """dlopen the in-tree CUDA library; raises (never falls back) when it is missing."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise FileNotFoundError(
            f"{p} is missing: build it with `python -m libultrahdr_dev_b200.build` "
            "(or __graft_entry__.build()). There is no CPU fallback.")
    lib = C.CDLL(p)
    for name, (res, args) in list(SYMBOLS.items()) + list(TEST_SYMBOLS.items()):
        f = getattr(lib, name)
        f.restype = res
        f.argtypes = args
    if lib.uhdrb200_abi_version() != 1:
        raise RuntimeError("libuhdr_b200.so ABI version mismatch")
    if path is None:
        _lib = lib
    return lib


def _check(lib, st):
    if st == abi.ERROR_INSUFFICIENT_RESOURCE:
        raise UltraHdrCudaError(lib.uhdrb200_last_error().decode() or "CUDA failure")
    return st


def _ref(x):
    return C.byref(x) if x is not None else None


class UltraHdr:
    """Mirror of `class ultrahdr::UltraHdr`'s protected hot-path members (ultrahdr.h:348-381).

    Descriptors are `abi.Image` / `abi.Metadata` (POD mirrors of the reference structs); plane
    pointers are HOST pointers for the plain methods and DEVICE pointers for the *Device /
    *Batch methods (which only enqueue on `stream`, a raw cudaStream_t integer).
    """

    def __init__(self, lib=None):
        self.lib = lib or load_library()

    # -- reference signatures, host planes --------------------------------------------------
    def generateGainMap(self, yuv420_image_ptr, p010_image_ptr, hdr_tf, metadata, dest, sdr_is_601=False):
        return _check(self.lib, self.lib.uhdrb200_generate_gainmap_host(
            _ref(yuv420_image_ptr), _ref(p010_image_ptr), hdr_tf, _ref(metadata), _ref(dest), 1 if sdr_is_601 else 0))

    def generateGainMapInto(self, yuv420_image_ptr, p010_image_ptr, hdr_tf, metadata, dest, map_buf, sdr_is_601=False):
        return _check(self.lib, self.lib.uhdrb200_generate_gainmap_host_into(
            _ref(yuv420_image_ptr), _ref(p010_image_ptr), hdr_tf, _ref(metadata), _ref(dest), 1 if sdr_is_601 else 0,
            map_buf.ctypes.data, map_buf.nbytes))

    def applyGainMap(self, yuv420_image_ptr, gainmap_image_ptr, metadata, output_format, max_display_boost, dest):
        return _check(self.lib, self.lib.uhdrb200_apply_gainmap_host(
            _ref(yuv420_image_ptr), _ref(gainmap_image_ptr), _ref(metadata), output_format, max_display_boost, _ref(dest)))

    def toneMap(self, src, dest):
        return _check(self.lib, self.lib.uhdrb200_tonemap_host(_ref(src), _ref(dest)))

    # -- device-resident, asynchronous ------------------------------------------------------
    def generateGainMapDevice(self, yuv420, p010, hdr_tf, metadata, dest, sdr_is_601=False, stream=0):
        return _check(self.lib, self.lib.uhdrb200_generate_gainmap_dev(
            _ref(yuv420), _ref(p010), hdr_tf, _ref(metadata), _ref(dest), 1 if sdr_is_601 else 0, stream))

    def applyGainMapDevice(self, yuv420, gainmap, metadata, output_format, max_display_boost, dest, stream=0):
        return _check(self.lib, self.lib.uhdrb200_apply_gainmap_dev(
            _ref(yuv420), _ref(gainmap), _ref(metadata), output_format, max_display_boost, _ref(dest), stream))

    def toneMapDevice(self, src, dest, stream=0):
        return _check(self.lib, self.lib.uhdrb200_tonemap_dev(_ref(src), _ref(dest), stream))

    # -- batched: ctypes arrays of descriptors, one launch ------------------------------------
    def generateGainMapBatch(self, n, yuv420s, p010s, hdr_tf, metadatas, dests, sdr_is_601=False, stream=0):
        return _check(self.lib, self.lib.uhdrb200_generate_gainmap_batch_dev(
            n, yuv420s, p010s, hdr_tf, metadatas, dests, 1 if sdr_is_601 else 0, stream))

    def applyGainMapBatch(self, n, yuv420s, gainmaps, metadata, output_format, max_display_boost, dests, stream=0):
        return _check(self.lib, self.lib.uhdrb200_apply_gainmap_batch_dev(
            n, yuv420s, gainmaps, _ref(metadata), output_format, max_display_boost, dests, stream))

    def toneMapBatch(self, n, srcs, dests, stream=0):
        return _check(self.lib, self.lib.uhdrb200_tonemap_batch_dev(n, srcs, dests, stream))


class JpegR(UltraHdr):
    """`JpegR : public UltraHdr` (jpegr.h:59) — the pixel steps on the path: convertYuv and the
    API-0 front half (toneMap + generateGainMap, jpegr.cpp:167-225) as one call."""

    def toneMapGenerateGainMapInto(self, p010, hdr_tf, sdr_dest, metadata, map_dest, map_buf, sdr_to_601=False):
        return _check(self.lib, self.lib.uhdrb200_tonemap_generate_gainmap_host_into(
            _ref(p010), hdr_tf, 1 if sdr_to_601 else 0, _ref(sdr_dest), _ref(metadata), _ref(map_dest),
            map_buf.ctypes.data, map_buf.nbytes))

    def toneMapGenerateGainMapDevice(self, p010, hdr_tf, sdr_dest, metadata, map_dest, stream=0, sdr_to_601=False):
        return _check(self.lib, self.lib.uhdrb200_tonemap_generate_gainmap_dev(
            _ref(p010), hdr_tf, 1 if sdr_to_601 else 0, _ref(sdr_dest), _ref(metadata), _ref(map_dest), stream))

    def toneMapGenerateGainMapBatch(self, n, p010s, hdr_tf, sdr_dests, metadatas, map_dests, stream=0, sdr_to_601=False):
        return _check(self.lib, self.lib.uhdrb200_tonemap_generate_gainmap_batch_dev(
            n, p010s, hdr_tf, 1 if sdr_to_601 else 0, sdr_dests, metadatas, map_dests, stream))

    def convertYuv(self, image, src_encoding, dest_encoding):
        return _check(self.lib, self.lib.uhdrb200_convert_yuv_host(_ref(image), src_encoding, dest_encoding))

    def convertYuvDevice(self, image, src_encoding, dest_encoding, stream=0):
        return _check(self.lib, self.lib.uhdrb200_convert_yuv_dev(_ref(image), src_encoding, dest_encoding, stream))

    # -- row N3: SDR output pack (examples/ultrahdr_app.cpp:682-734) and copy-with-restride (jpegr.cpp:315-359)
    def yuv420ToRgba8888(self, yuv420, dest):
        return _check(self.lib, self.lib.uhdrb200_yuv420_to_rgba8888_host(_ref(yuv420), _ref(dest)))

    def yuv420ToRgba8888Device(self, yuv420, dest, stream=0):
        return _check(self.lib, self.lib.uhdrb200_yuv420_to_rgba8888_dev(_ref(yuv420), _ref(dest), stream))

    def yuv420ToRgba8888Batch(self, n, yuv420s, dests, stream=0):
        return _check(self.lib, self.lib.uhdrb200_yuv420_to_rgba8888_batch_dev(n, yuv420s, dests, stream))

    def copyYuv420(self, src, dest):
        return _check(self.lib, self.lib.uhdrb200_copy_yuv420_host(_ref(src), _ref(dest)))

    def copyYuv420Device(self, src, dest, stream=0):
        return _check(self.lib, self.lib.uhdrb200_copy_yuv420_dev(_ref(src), _ref(dest), stream))


def take_map(dest):
    """Copy a malloc'ed gain map returned by generateGainMap into numpy and free it."""
    import numpy as np

    n = dest.width * dest.height
    arr = np.ctypeslib.as_array(C.cast(dest.data, C.POINTER(C.c_uint8)), shape=(max(n, 1),))[:n].copy()
    load_library().uhdrb200_free(dest.data)
    dest.data = None
    return arr.reshape(dest.height, dest.width)
