#!/bin/bash
# ncu --set full captures of the two C4 throughput-path kernels (run on the GPU box AFTER the plain command exited 0)
set -e
B="python bench.py --steps 2 --warmup 1 --no-latency --no-cpu --no-configs"
$B > gpurun_out/plain_c4.log 2>&1
N="ncu --set full --clock-control none --import-source on -f"
$N -k regex:forest_rank_keys_kernel -s 20 -c 1 -o gpurun_out/ncu_c4_keys_r02 $B > gpurun_out/ncu_c4a.log 2>&1
$N -k regex:forest_compact_kernel -s 20 -c 1 -o gpurun_out/ncu_c4_walk_r02 $B > gpurun_out/ncu_c4b.log 2>&1
for k in keys walk; do python tools/ncu_summary.py gpurun_out/ncu_c4_${k}_r02.ncu-rep "tools/ncu_c4.sh" > gpurun_out/ncu_c4_${k}_r02.txt; ncu -i gpurun_out/ncu_c4_${k}_r02.ncu-rep --page source --csv --print-source sass > gpurun_out/ncu_c4_${k}_src.csv; done
rm -f gpurun_out/ncu_c4_keys_r02.ncu-rep gpurun_out/ncu_c4_walk_r02.ncu-rep
ls -la gpurun_out/ncu_c4_*
