#!/bin/bash
# ncu evidence of the final tree (names r02z_*): (1) launch list of the bench command restricted to the gain-map kernels
# (the 256-slot set-up launches >1500 torch kernels first), (2) one --set full capture of every gain-map kernel
# instantiation, (3) one --set full capture of the JPEG codec kernels on a 4K quality-90 decode + encode.
cd "$(dirname "$0")/.." || exit 1
K='regex:generate_kernel|apply_q_kernel|apply_qs_kernel|apply_fast_kernel|apply_kernel|apply_pipe_kernel|tonemap_kernel|convert_yuv|rgba8888|copy_yuv420|jpeg_idct'
A="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
B="python bench.py --steps 2 --warmup 3 --batch 64 --distinct 4 --no-cpu-baseline --no-e2e --profile-once"
$A > gpurun_out/r02z_plain_a.json 2> gpurun_out/r02z_plain_a.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/r02z_launches.csv $A > gpurun_out/r02z_ncu_a.log 2>&1
$B > gpurun_out/r02z_plain_b.json 2> gpurun_out/r02z_plain_b.err &&
ncu --set full --clock-control none --import-source on -k "$K" -s 8 -c 40 -o gpurun_out/r02z_full $B > gpurun_out/r02z_ncu_b.log 2>&1
python scripts/ncu_show.py gpurun_out/r02z_full.ncu-rep > gpurun_out/r02z_ncu_details.txt 2>&1
python scripts/ncu_traffic.py gpurun_out/r02z_full.ncu-rep $((64*3840*2160)) gpurun_out/r02z_ncu_full > gpurun_out/r02z_traffic.log 2>&1
[ "$(stat -c %s gpurun_out/r02z_full.ncu-rep)" -gt 30000000 ] && rm -f gpurun_out/r02z_full.ncu-rep
cat > /tmp/codec_once.py <<'PY'
import ctypes as C, io, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from PIL import Image
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
for _ in range(2): assert lib.uhdrb200_jpeg_encode_dev(C.byref(img), 90, None, 0, out.ctypes.data, out.size, C.byref(n), None) == 0
buf = (C.c_ubyte * n.value).from_buffer_copy(out[: n.value].tobytes())
d2 = torch.zeros(W * H * 3 // 2, dtype=torch.uint8, device="cuda"); im2 = abi.Image(); im2.data = d2.data_ptr()
for _ in range(2): assert lib.uhdrb200_jpeg_decode_dev(buf, n.value, C.byref(im2), None) == 0
print("codec ok", n.value)
PY
python /tmp/codec_once.py > gpurun_out/r02z_codec_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:huff_|fdct_|dc_|scan_|jpeg_idct' -s 14 -c 16 -o gpurun_out/r02z_codec python /tmp/codec_once.py > gpurun_out/r02z_ncu_codec.log 2>&1
python scripts/ncu_show.py gpurun_out/r02z_codec.ncu-rep > gpurun_out/r02z_ncu_codec_details.txt 2>&1
[ "$(stat -c %s gpurun_out/r02z_codec.ncu-rep)" -gt 20000000 ] && rm -f gpurun_out/r02z_codec.ncu-rep
ls -la gpurun_out/ | grep r02z
