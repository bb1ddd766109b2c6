#!/bin/bash
# one --set full capture of the IDCT kernel on the bench's 8160x6144 4:2:0 coefficient planes (after a plain run of the same program)
cd "$(dirname "$0")/.." || exit 1
TAG=${TAG:-r02zz}
cat > /tmp/idct_once.py <<'PY'
import sys
sys.path.insert(0, "/root/repo")
import torch
from libultrahdr_dev_b200 import abi, api
lib = api.JpegR().lib
dev = torch.device("cuda:0")
jw, jh = 8160, 6144
jn = jw * jh * 3 // 2
torch.manual_seed(0)
d_co = (torch.randint(-40, 40, (jn,), dtype=torch.int16, device=dev) * (torch.rand((jn,), device=dev) < 0.12).to(torch.int16))
d_pl = torch.zeros((jn,), dtype=torch.uint8, device=dev)
d_q = torch.randint(1, 32, (64,), dtype=torch.int16, device=dev)
jc = (abi.JpegComponent * 3)()
for k, (off, cw, ch) in enumerate(((0, jw, jh), (jw * jh, jw // 2, jh // 2), (jw * jh * 5 // 4, jw // 2, jh // 2))):
    jc[k].coeffs, jc[k].quant = d_co.data_ptr() + 2 * off, d_q.data_ptr()
    jc[k].blocks_w, jc[k].blocks_h = cw // 8, ch // 8
    jc[k].plane, jc[k].width, jc[k].height, jc[k].stride = d_pl.data_ptr() + off, cw, ch, cw
torch.cuda.synchronize()
for _ in range(3): assert lib.uhdrb200_jpeg_idct_dev(3, jc, 0) == 0
torch.cuda.synchronize()
print("idct ok", int(d_pl.sum()))
PY
python /tmp/idct_once.py > gpurun_out/${TAG}_idct_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:jpeg_idct' -s 2 -c 1 -o gpurun_out/${TAG}_idct python /tmp/idct_once.py > gpurun_out/${TAG}_ncu_idct.log 2>&1
python scripts/ncu_show.py gpurun_out/${TAG}_idct.ncu-rep > gpurun_out/${TAG}_ncu_idct_details.txt 2>&1
ncu -i gpurun_out/${TAG}_idct.ncu-rep --page source --csv > gpurun_out/${TAG}_idct_source.csv 2>/dev/null
cat gpurun_out/${TAG}_ncu_idct_details.txt
