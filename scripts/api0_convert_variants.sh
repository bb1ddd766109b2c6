#!/bin/bash
# A/B of the fused API-0 + convertYuv kernel's occupancy bound (UHDR_API0_CONVERT_MIN_BLOCKS): builds on the box
cd "$(dirname "$0")/.." || exit 1
cat > /tmp/api0.py <<'PY'
import sys, json
sys.path.insert(0, "/root/repo")
import numpy as np, torch, ctypes as C
from libultrahdr_dev_b200 import abi, api, synth
uh = api.JpegR(); W, H, B = 3840, 2160, 64
dev = torch.device("cuda", 0)
pairs = [synth.natural_pair(W, H, i, abi.TF_HLG, 2, 2) for i in range(4)]
d_hdr = torch.stack([torch.from_numpy(pairs[i % 4][1].view(np.int16)).roll(128 * i) for i in range(B)]).to(dev)
d_sdr = torch.zeros((B, W * H * 3 // 2), dtype=torch.uint8, device=dev)
d_map = torch.zeros((B, W * H // 16), dtype=torch.uint8, device=dev)
Img = abi.Image * B; hdrs, sdrs, maps = Img(), Img(), Img(); mds = (abi.Metadata * B)()
for i in range(B):
    hb = d_hdr[i].data_ptr(); hdrs[i].data, hdrs[i].chroma_data = hb, hb + 2 * W * H
    hdrs[i].width, hdrs[i].height, hdrs[i].colorGamut, hdrs[i].luma_stride, hdrs[i].chroma_stride, hdrs[i].pixelFormat = W, H, 2, W, W, abi.PIX_FMT_P010
    sb = d_sdr[i].data_ptr(); sdrs[i].data, sdrs[i].chroma_data = sb, sb + W * H
    sdrs[i].width, sdrs[i].height, sdrs[i].colorGamut, sdrs[i].luma_stride, sdrs[i].chroma_stride, sdrs[i].pixelFormat = W, H, -1, W, W // 2, abi.PIX_FMT_YUV420
    maps[i].data = d_map[i].data_ptr()
st = torch.cuda.current_stream().cuda_stream
res = {}
for name, to601 in (("fused", False), ("fused_convert", True)):
    for _ in range(3): assert uh.toneMapGenerateGainMapBatch(B, hdrs, abi.TF_HLG, sdrs, mds, maps, st, sdr_to_601=to601) == 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): uh.toneMapGenerateGainMapBatch(B, hdrs, abi.TF_HLG, sdrs, mds, maps, st, sdr_to_601=to601)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    res[name] = {"gpix_s": round(B * W * H / ms / 1e6), "frac": round(B * W * H * 4.5625 / (ms * 1e-3) / 1e9 / 6540.8, 3)}
print(json.dumps(res))
PY
for mb in ${MB_LIST:-1 4 5 6}; do
  UHDR_NVCC_EXTRA="-D${MB_MACRO:-UHDR_API0_CONVERT_MIN_BLOCKS}=$mb" python -m libultrahdr_dev_b200.build --force > /dev/null 2>&1 || { echo build failed; continue; }
  echo "min_blocks=$mb $(python /tmp/api0.py 2>&1 | tail -1)"
done
