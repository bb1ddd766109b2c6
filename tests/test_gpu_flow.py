"""The pixel side of JpegR::encodeJPEGR API-0 followed by JpegR::decodeJPEGR (lib/src/jpegr.cpp:167-261, 693-801), every
step on the device and nothing but compressed bytes crossing back: P010 -> toneMap + generateGainMap + convertYuv (one
kernel) -> JPEG encode of the SDR image and of the gain map -> JPEG decode of both -> applyGainMap. Each hand-over is
checked against the CPU side: the oracle for the pixel math, libjpeg-turbo (through Pillow) for the two codecs."""
import ctypes as C

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth

import parity

pytestmark = pytest.mark.gpu
FLT_MAX = 3.4028234663852886e38


def test_api0_encode_then_decode_flow_stays_on_device(uh, orc):
    pytest.importorskip("PIL.Image")
    import torch as t

    from test_gpu_jpeg_decode import libjpeg_planes
    from test_gpu_jpeg_encode import pillow_encode

    lib = uh.lib
    w, h = 1280, 720
    _, hdr_b = synth.natural_pair(w, h, 31, abi.TF_HLG, 2, 2)
    st = t.cuda.current_stream().cuda_stream
    d_hdr = t.from_numpy(hdr_b.view(np.int16)).cuda()
    d_sdr = t.zeros(w * h * 3 // 2, dtype=t.uint8, device="cuda")
    d_map = t.zeros(w * h // 16, dtype=t.uint8, device="cuda")
    hdr = abi.host_p010(hdr_b, w, h, abi.COLORGAMUT_BT2100)
    hdr.data, hdr.chroma_data = d_hdr.data_ptr(), d_hdr.data_ptr() + 2 * w * h
    sdr, mp, md = abi.Image(), abi.Image(), abi.Metadata()
    sdr.data, sdr.chroma_data = d_sdr.data_ptr(), d_sdr.data_ptr() + w * h
    sdr.width, sdr.height, sdr.luma_stride, sdr.chroma_stride, sdr.pixelFormat, sdr.colorGamut = w, h, w, w // 2, abi.PIX_FMT_YUV420, -1
    mp.data = d_map.data_ptr()
    # ---- encode side: jpegr.cpp:199-225 in one kernel, then the two compressImage calls (:234-239, :1037-1050)
    assert uh.toneMapGenerateGainMapDevice(hdr, abi.TF_HLG, sdr, md, mp, st, True) == 0
    files = []
    for img, q in ((sdr, 95), (mp, 85)):
        out = np.zeros(w * h * 3, np.uint8)
        n = C.c_size_t(0)
        assert lib.uhdrb200_jpeg_encode_dev(C.byref(img), q, None, 0, out.ctypes.data, out.size, C.byref(n), st) == 0, lib.uhdrb200_last_error()
        files.append(bytes(out[: n.value]))
    sdr_planes, map_plane = d_sdr.cpu().numpy(), d_map.cpu().numpy().reshape(h // 4, w // 4)
    # the pixel math against the oracle chain (toneMap, generate on the un-converted image, convertYuv -> BT.601)
    so = np.zeros_like(sdr_planes)
    sod = abi.host_yuv420(so, w, h, -1)
    hh = abi.host_p010(hdr_b, w, h, abi.COLORGAMUT_BT2100)
    assert orc.tonemap(hh, sod) == 0
    _, want_map, md_o, _ = orc.generate_gainmap(sod, hh, abi.TF_HLG)
    assert orc.convert_yuv(sod, abi.COLORGAMUT_BT2100, abi.COLORGAMUT_P3) == 0
    assert np.array_equal(sdr_planes, so) and md.as_dict() == md_o.as_dict()
    parity.compare_bytes(map_plane, want_map, "flow: gain map", min_exact=0.999)
    # the two files against libjpeg-turbo fed with the same planes
    y = sdr_planes[: w * h].reshape(h, w)
    u = sdr_planes[w * h: w * h * 5 // 4].reshape(h // 2, w // 2)
    v = sdr_planes[w * h * 5 // 4:].reshape(h // 2, w // 2)
    assert files[0] == pillow_encode([y, u, v], 95)
    assert (h // 4) % 8 != 0 or files[1] == pillow_encode([map_plane], 85)  # the 320 x 180 map has a partial block row: covered by the zero-padding test
    print(f"flow: SDR JPEG {len(files[0])} bytes, gain-map JPEG {len(files[1])} bytes for {w * h * 3} bytes of P010")
    # ---- decode side: jpegr.cpp:693 / :732 (two decodes), :801 (applyGainMap)
    d_sdr2 = t.zeros(w * h * 3 // 2, dtype=t.uint8, device="cuda")
    d_map2 = t.zeros(w * h // 16, dtype=t.uint8, device="cuda")
    d_out = t.zeros(w * h, dtype=t.int32, device="cuda")
    sdr2, mp2, dst = abi.Image(), abi.Image(), abi.Image()
    sdr2.data, mp2.data, dst.data = d_sdr2.data_ptr(), d_map2.data_ptr(), d_out.data_ptr()
    for data, img in ((files[0], sdr2), (files[1], mp2)):
        buf = (C.c_ubyte * len(data)).from_buffer_copy(data)
        assert lib.uhdrb200_jpeg_decode_dev(buf, len(data), C.byref(img), st) == 0, lib.uhdrb200_last_error()
    sdr2.colorGamut = abi.COLORGAMUT_BT2100
    assert lib.uhdrb200_apply_gainmap_dev(C.byref(sdr2), C.byref(mp2), C.byref(md), abi.OUTPUT_HDR_HLG, FLT_MAX, C.byref(dst), st) == 0
    t.cuda.synchronize()
    ref_sdr = np.concatenate([p.ravel() for p in libjpeg_planes(files[0])])
    ref_map = libjpeg_planes(files[1])[0]
    assert np.array_equal(d_sdr2.cpu().numpy(), ref_sdr) and np.array_equal(d_map2.cpu().numpy(), ref_map.ravel())
    _, want, _ = orc.apply_gainmap(abi.host_yuv420(ref_sdr, w, h, abi.COLORGAMUT_BT2100), abi.host_map(ref_map, w // 4, h // 4), md,
                                   abi.OUTPUT_HDR_HLG, FLT_MAX)
    parity.compare_output(abi.OUTPUT_HDR_HLG, d_out.cpu().numpy().view(np.uint8), want, w * h, "flow: recovered HDR", min_exact=0.995)
