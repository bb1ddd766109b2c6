"""apply-only timing on flat / smooth (noise-free) / natural / uniform-random images."""
import sys, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from libultrahdr_dev_b200 import abi, api, synth
lib = api.load_library(); uh = api.JpegR(lib)
W, H, B = 3840, 2160, 32
fmt = int(sys.argv[1]) if len(sys.argv) > 1 else 3
dev = torch.device("cuda", 0)
md = abi.default_metadata(1000.0 / 203.0)
def run(name, sdr, mp):
    d_sdr = torch.empty((B, sdr.size), dtype=torch.uint8, device=dev); d_sdr[:] = torch.from_numpy(sdr).to(dev)
    d_map = torch.empty((B, W * H // 16), dtype=torch.uint8, device=dev); d_map[:] = torch.from_numpy(mp.ravel()).to(dev)
    d_out = torch.zeros((B, W * H * abi.out_bytes_per_pixel(fmt)), dtype=torch.uint8, device=dev)
    Img = abi.Image * B; sdrs, mps, outs = Img(), Img(), Img()
    for i in range(B):
        b = d_sdr[i].data_ptr(); sdrs[i].data, sdrs[i].chroma_data = b, b + W * H
        sdrs[i].width, sdrs[i].height, sdrs[i].colorGamut = W, H, 0
        sdrs[i].luma_stride, sdrs[i].chroma_stride, sdrs[i].pixelFormat = W, W // 2, abi.PIX_FMT_YUV420
        mps[i].data, mps[i].width, mps[i].height, mps[i].luma_stride = d_map[i].data_ptr(), W // 4, H // 4, W // 4
        outs[i].data = d_out[i].data_ptr()
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(3): uh.applyGainMapBatch(B, sdrs, mps, md, fmt, 3.4e38, outs, st)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): uh.applyGainMapBatch(B, sdrs, mps, md, fmt, 3.4e38, outs, st)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"{name:10s} fmt {fmt}: {B*W*H/ms/1e6:8.1f} Gpix/s  {B*W*H*(1.5625+abi.out_bytes_per_pixel(fmt))/ms/1e6:7.0f} GB/s")
flat = np.full(abi.yuv420_size(W, H), 120, dtype=np.uint8)
run("flat", flat, np.full((H // 4, W // 4), 128, dtype=np.uint8))
nat = synth.natural_pair(W, H, 1)[0]
# noise-free smooth: blur the natural image's luma heavily (box filter by down/up-sampling)
y = nat[:W * H].reshape(H, W); ys = y.reshape(H // 8, 8, W // 8, 8).mean(axis=(1, 3)); ysm = np.repeat(np.repeat(ys, 8, 0), 8, 1).astype(np.uint8)
smooth = nat.copy(); smooth[:W * H] = ysm.ravel()
run("smooth8x8", smooth, synth.natural_map(W // 4, H // 4, 2))
run("natural", nat, synth.natural_map(W // 4, H // 4, 2))
run("uniform", synth.uniform_pair(W, H, 1)[0], np.random.Generator(np.random.PCG64(1)).integers(0, 256, (H // 4, W // 4), dtype=np.uint8))
