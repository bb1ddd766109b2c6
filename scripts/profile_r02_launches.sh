#!/bin/bash
# ncu launch list of the bench command on the last tree (names ${TAG}_*), this library's kernels only, after a plain run of
# the same command; then one --set full capture of the JPEG encoder's kernels (scripts/ncu_encoder.sh).
cd "$(dirname "$0")/.." || exit 1
TAG=${TAG:-r02zzz}
K='regex:generate_kernel|apply_q_kernel|apply_qs_kernel|apply_fast_kernel|apply_kernel|apply_pipe_kernel|tonemap_kernel|convert_yuv|rgba8888|copy_yuv420|jpeg_idct|fdct_|huff_|dc_|scan_'
A="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
$A > gpurun_out/${TAG}_bench_profiled_command.json 2> gpurun_out/${TAG}_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" -c 700 --csv --log-file gpurun_out/${TAG}_launches.csv $A > gpurun_out/${TAG}_ncu_a.log 2>&1
wc -l gpurun_out/${TAG}_launches.csv
bash scripts/ncu_encoder.sh | tail -12
