cd /root/repo
cat > /tmp/one.py <<'PY'
import ctypes as C, io, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from PIL import Image
from libultrahdr_dev_b200 import abi, api, synth
uh = api.JpegR(); lib = uh.lib
W, H = 3840, 2160
sdr, _ = synth.natural_pair(W, H, 0)
y = sdr[: W * H].reshape(H, W)
u = sdr[W * H: W * H * 5 // 4].reshape(H // 2, W // 2).repeat(2, 0).repeat(2, 1)
v = sdr[W * H * 5 // 4:].reshape(H // 2, W // 2).repeat(2, 0).repeat(2, 1)
rng = np.random.Generator(np.random.PCG64(0))
y = np.clip(y.astype(np.int16) + rng.integers(-6, 7, y.shape), 0, 255).astype(np.uint8)
b = io.BytesIO(); Image.fromarray(np.stack([y, u, v], -1), "YCbCr").save(b, "JPEG", quality=90, subsampling=2); d = b.getvalue()
buf = (C.c_ubyte * len(d)).from_buffer_copy(d)
d_yuv = torch.zeros(W * H * 3 // 2, dtype=torch.uint8, device="cuda")
img = abi.Image(); img.data = d_yuv.data_ptr()
for _ in range(2): assert lib.uhdrb200_jpeg_decode_dev(buf, len(d), C.byref(img), 0) == 0
PY
python /tmp/one.py && ncu --set full --clock-control none --import-source on -k regex:huff_ -s 4 -c 4 -o gpurun_out/r02y_huff python /tmp/one.py > gpurun_out/r02y_ncu.log 2>&1
ncu -i gpurun_out/r02y_huff.ncu-rep --page details --csv 2>/dev/null | python - <<'PY'
import csv, sys
rows = list(csv.DictReader(sys.stdin))
want = ("Duration", "Executed Ipc Active", "Registers Per Thread", "Achieved Occupancy", "Issue Slots Busy", "Warp Cycles Per Issued Instruction", "Executed Instructions", "L1/TEX Hit Rate", "Avg. Active Threads Per Warp", "Avg. Not Predicated Off Threads Per Warp", "Elapsed Cycles", "No Eligible", "One or More Eligible")
for r in rows:
    if r["Metric Name"] in want:
        print(r["ID"], r["Kernel Name"][:40], "|", r["Metric Name"], "=", r["Metric Value"], r["Metric Unit"])
PY
ls -la gpurun_out/r02y_huff.ncu-rep
