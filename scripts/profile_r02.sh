#!/bin/bash
# ncu evidence for the round: (1) launch list of the bench command, (2) one --set full capture of every kernel instantiation
cd "$(dirname "$0")/.." || exit 1
A="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
B="python bench.py --steps 2 --warmup 3 --batch 64 --distinct 4 --no-cpu-baseline --no-e2e --profile-once"
$A > gpurun_out/r02l_plain_a.json 2> gpurun_out/r02l_plain_a.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02l_launches.csv $A > gpurun_out/r02l_ncu_a.log 2>&1
$B > gpurun_out/r02l_plain_b.json 2> gpurun_out/r02l_plain_b.err &&
ncu --set full --clock-control none --import-source on -k regex:'apply_q_kernel|apply_qs_kernel|generate_kernel|tonemap_kernel|convert_yuv|rgba8888' -s 10 -c 40 -o gpurun_out/r02l_full $B > gpurun_out/r02l_ncu_b.log 2>&1
ls -la gpurun_out/ | grep r02l
