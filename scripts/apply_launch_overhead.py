"""Fixed cost per applyGainMap launch: per-launch GPU time t(n) of the batched entry point for n = 1, 2, 3, 4, 8, 16 4K
images (back-to-back launches on one stream, images rotated), so that t(n) = t0 + n * t1 separates the per-launch
overhead t0 (ramp, table staging, tail imbalance) from the steady-state cost per image t1."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from libultrahdr_dev_b200 import abi, api, synth

uh = api.JpegR(); W, H, B = 3840, 2160, 48
dev = torch.device("cuda", 0)
pairs = [synth.natural_pair(W, H, i) for i in range(4)]
d_sdr = torch.stack([torch.from_numpy(pairs[i % 4][0]).to(dev).roll(64 * i) for i in range(B)])
d_map = torch.stack([torch.from_numpy(synth.natural_map(W // 4, H // 4, i % 4)).to(dev).reshape(-1).roll(16 * i) for i in range(B)])
d_out = torch.zeros((B, W * H), dtype=torch.int32, device=dev)
Img = abi.Image * B; sdrs, maps, outs = Img(), Img(), Img()
for i in range(B):
    b = d_sdr[i].data_ptr(); sdrs[i].data, sdrs[i].chroma_data = b, b + W * H
    sdrs[i].width, sdrs[i].height, sdrs[i].colorGamut, sdrs[i].luma_stride, sdrs[i].chroma_stride, sdrs[i].pixelFormat = W, H, 0, W, W // 2, abi.PIX_FMT_YUV420
    maps[i].data, maps[i].width, maps[i].height, maps[i].luma_stride, maps[i].pixelFormat = d_map[i].data_ptr(), W // 4, H // 4, W // 4, abi.PIX_FMT_MONOCHROME
    outs[i].data = d_out[i].data_ptr()
st = torch.cuda.current_stream().cuda_stream; md = abi.default_metadata(1000.0 / 203.0)
import ctypes as C
def sub(arr, off, n):
    return C.cast(C.byref(arr, off * C.sizeof(abi.Image)), C.POINTER(abi.Image))
res = {}
for n in (1, 2, 3, 4, 8, 16):
    offs = list(range(0, B - n + 1, n))
    def go(k):
        o = offs[k % len(offs)]
        return uh.lib.uhdrb200_apply_gainmap_batch_dev(n, sub(sdrs, o, n), sub(maps, o, n), C.byref(md), abi.OUTPUT_HDR_HLG, 3.4e38, sub(outs, o, n), st)
    for k in range(6): assert go(k) == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    iters = 96
    torch.cuda.synchronize(); e0.record()
    for k in range(iters): go(k)
    e1.record(); torch.cuda.synchronize()
    res[n] = round(e0.elapsed_time(e1) / iters * 1e3, 2)
ns = np.array(list(res.keys()), dtype=np.float64); ts = np.array(list(res.values()))
t1, t0 = np.polyfit(ns, ts, 1)
print(json.dumps({"us_per_launch": res, "fit_t0_us": round(float(t0), 2), "fit_t1_us_per_image": round(float(t1), 2), "env": {k: v for k, v in os.environ.items() if k.startswith("UHDRB200")}}))
