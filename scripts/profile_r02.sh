#!/bin/bash
# ncu evidence for the round: (1) launch list of the bench command (this library's kernels only: the set-up of the 256
# device slots launches >1500 torch kernels first), (2) one --set full capture of every kernel instantiation
cd "$(dirname "$0")/.." || exit 1
K='regex:generate_kernel|apply_q_kernel|apply_qs_kernel|apply_fast_kernel|apply_kernel|apply_pipe_kernel|tonemap_kernel|convert_yuv|rgba8888|copy_yuv420|jpeg_idct'
A="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
B="python bench.py --steps 2 --warmup 3 --batch 64 --distinct 4 --no-cpu-baseline --no-e2e --profile-once"
$A > gpurun_out/r02s_plain_a.json 2> gpurun_out/r02s_plain_a.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 400 --csv --log-file gpurun_out/r02s_launches.csv $A > gpurun_out/r02s_ncu_a.log 2>&1
$B > gpurun_out/r02s_plain_b.json 2> gpurun_out/r02s_plain_b.err &&
ncu --set full --clock-control none --import-source on -k "$K" -s 8 -c 36 -o gpurun_out/r02s_full $B > gpurun_out/r02s_ncu_b.log 2>&1
ls -la gpurun_out/ | grep r02s
# summaries are made on the box (gpurun_out/ is merged back only below 64 MiB): keep the report itself only if it is small
python scripts/ncu_show.py gpurun_out/r02s_full.ncu-rep > gpurun_out/r02s_ncu_details.txt 2>&1
python scripts/ncu_traffic.py gpurun_out/r02s_full.ncu-rep $((64*3840*2160)) gpurun_out/r02s_ncu_full > gpurun_out/r02s_traffic.log 2>&1
[ "$(stat -c %s gpurun_out/r02s_full.ncu-rep)" -gt 45000000 ] && rm -f gpurun_out/r02s_full.ncu-rep
ls -la gpurun_out/ | grep r02s
