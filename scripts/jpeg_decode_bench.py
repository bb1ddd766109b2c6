"""Device JPEG decode (uhdrb200_jpeg_decode_dev: host parse + byte-stuffing removal, upload of the compressed bytes,
parallel Huffman decode, IDCT) of 4K 4:2:0 images at a few qualities, wall clock per call on one host thread, next to
libjpeg-turbo (through Pillow) on one core."""
import ctypes as C, io, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from PIL import Image
from libultrahdr_dev_b200 import abi, api, synth

uh = api.JpegR(); lib = uh.lib
W, H = 3840, 2160
dev = torch.device("cuda", 0)


def ycc_image(i):
    sdr, _ = synth.natural_pair(W, H, i)
    y = sdr[: W * H].reshape(H, W)
    u = sdr[W * H: W * H * 5 // 4].reshape(H // 2, W // 2).repeat(2, 0).repeat(2, 1)
    v = sdr[W * H * 5 // 4:].reshape(H // 2, W // 2).repeat(2, 0).repeat(2, 1)
    rng = np.random.Generator(np.random.PCG64(i))
    y = np.clip(y.astype(np.int16) + rng.integers(-6, 7, y.shape), 0, 255).astype(np.uint8)  # sensor-like texture
    return Image.fromarray(np.stack([y, u, v], -1), "YCbCr")


imgs = [ycc_image(i) for i in range(4)]
d_yuv = torch.zeros((4, W * H * 3 // 2), dtype=torch.uint8, device=dev)
st = torch.cuda.current_stream().cuda_stream
out = []
for q, extra in ((75, {}), (90, {}), (95, {}), (100, {}), (90, {"restart_marker_rows": 1}), (100, {"restart_marker_rows": 1}), (90, {"restart_marker_blocks": 16})):
    datas = []
    for im in imgs:
        b = io.BytesIO(); im.save(b, "JPEG", quality=q, subsampling=2, **extra); datas.append(b.getvalue())
    bufs = [(C.c_ubyte * len(d)).from_buffer_copy(d) for d in datas]
    dst = [abi.Image() for _ in range(4)]
    for k in range(4): dst[k].data = d_yuv[k].data_ptr()
    n0 = lib.uhdrb200_launch_count()
    for k in range(4): assert lib.uhdrb200_jpeg_decode_dev(bufs[k], len(datas[k]), C.byref(dst[k]), st) == 0, lib.uhdrb200_last_error()
    passes = (lib.uhdrb200_launch_count() - n0) / 4 - 7
    iters = 16
    t0 = time.perf_counter()
    for k in range(iters): lib.uhdrb200_jpeg_decode_dev(bufs[k % 4], len(datas[k % 4]), C.byref(dst[k % 4]), st)
    dt = (time.perf_counter() - t0) / iters
    t0 = time.perf_counter()
    for k in range(iters): lib.uhdrb200_jpeg_get_info(bufs[k % 4], len(datas[k % 4]), None, None, None)
    dt_parse = (time.perf_counter() - t0) / iters
    t0 = time.perf_counter()
    for k in range(4):
        im = Image.open(io.BytesIO(datas[k])); im.draft("YCbCr", (W, H)); im.load()
    dt_cpu = (time.perf_counter() - t0) / 4
    out.append({"quality": q, "restart": extra or None, "jpeg_bytes": int(np.mean([len(d) for d in datas])), "sync_passes": passes, "decode_dev_ms": round(dt * 1e3, 3),
                "mpix_s": round(W * H / dt / 1e6), "parse_only_ms": round(dt_parse * 1e3, 3), "libjpeg_turbo_pillow_1core_ms": round(dt_cpu * 1e3, 2),
                "libjpeg_turbo_mpix_s": round(W * H / dt_cpu / 1e6)})
    print(json.dumps(out[-1]), flush=True)

# ---- encode: 4K planes already in device memory -> JFIF bytes in host memory, against libjpeg-turbo on one core
planes = []
for i in range(4):
    sdr, _ = synth.natural_pair(W, H, i)
    rng = np.random.Generator(np.random.PCG64(i))
    sdr = sdr.copy(); sdr[: W * H] = np.clip(sdr[: W * H].astype(np.int16) + rng.integers(-6, 7, W * H), 0, 255).astype(np.uint8)
    planes.append(sdr)
d_pl = torch.stack([torch.from_numpy(p) for p in planes]).to(dev)
src = [abi.Image() for _ in range(4)]
for k in range(4):
    b = d_pl[k].data_ptr()
    src[k].data, src[k].chroma_data = b, b + W * H
    src[k].width, src[k].height, src[k].luma_stride, src[k].chroma_stride, src[k].pixelFormat = W, H, W, W // 2, abi.PIX_FMT_YUV420
outb = np.zeros(W * H * 3, np.uint8); nb = C.c_size_t(0)
for q in (75, 90, 95, 100):
    for k in range(4): assert lib.uhdrb200_jpeg_encode_dev(C.byref(src[k]), q, None, 0, outb.ctypes.data, outb.size, C.byref(nb), st) == 0, lib.uhdrb200_last_error()
    iters = 16
    t0 = time.perf_counter()
    for k in range(iters): lib.uhdrb200_jpeg_encode_dev(C.byref(src[k % 4]), q, None, 0, outb.ctypes.data, outb.size, C.byref(nb), st)
    dt = (time.perf_counter() - t0) / iters
    t0 = time.perf_counter()
    for k in range(2):
        bio = io.BytesIO(); imgs[k].save(bio, "JPEG", quality=q, subsampling=2)
    dt_cpu = (time.perf_counter() - t0) / 2
    print(json.dumps({"encode_quality": q, "jpeg_bytes": int(nb.value), "encode_dev_ms": round(dt * 1e3, 3), "mpix_s": round(W * H / dt / 1e6),
                      "libjpeg_turbo_pillow_1core_ms": round(dt_cpu * 1e3, 2), "libjpeg_turbo_mpix_s": round(W * H / dt_cpu / 1e6)}), flush=True)
