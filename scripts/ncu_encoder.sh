#!/bin/bash
# one --set full capture of the JPEG encoder's kernels on a 4K quality-90 encode (after a plain run of the same program)
cd "$(dirname "$0")/.." || exit 1
cat > /tmp/enc_once.py <<'PY'
import ctypes as C, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from libultrahdr_dev_b200 import abi, api, synth
lib = api.JpegR().lib
W, H = 3840, 2160
sdr, _ = synth.natural_pair(W, H, 0)
rng = np.random.Generator(np.random.PCG64(0))
sdr = sdr.copy(); sdr[: W * H] = np.clip(sdr[: W * H].astype(np.int16) + rng.integers(-6, 7, W * H), 0, 255).astype(np.uint8)
d = torch.from_numpy(sdr).cuda()
img = abi.Image(); img.data, img.chroma_data = d.data_ptr(), d.data_ptr() + W * H
img.width, img.height, img.luma_stride, img.chroma_stride, img.pixelFormat = W, H, W, W // 2, abi.PIX_FMT_YUV420
out = np.zeros(W * H * 3, np.uint8); n = C.c_size_t(0)
for _ in range(3): assert lib.uhdrb200_jpeg_encode_dev(C.byref(img), 90, None, 0, out.ctypes.data, out.size, C.byref(n), None) == 0
print("encode ok", n.value)
PY
python /tmp/enc_once.py > gpurun_out/r02z_enc_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:fdct_|huff_size|huff_emit|scan_' -s 12 -c 6 -o gpurun_out/r02z_enc python /tmp/enc_once.py > gpurun_out/r02z_ncu_enc.log 2>&1
python scripts/ncu_show.py gpurun_out/r02z_enc.ncu-rep > gpurun_out/r02z_ncu_encoder_details.txt 2>&1
rm -f gpurun_out/r02z_enc.ncu-rep
grep -E "^==|time_duration|inst_executed.sum|warps_active" gpurun_out/r02z_ncu_encoder_details.txt | head -30
