#!/bin/bash
# ncu --set full captures of the kernels behind C2 / C3 / C5-MLP (run on the GPU box AFTER the plain commands exited 0):
#   bash tools/ncu_configs.sh        -> gpurun_out/ncu_{c2,c3,mlp_layer,mlp_slice,mlp_fp64}_r02.ncu-rep
set -e
run() { python tools/bench_configs.py --only "$1" > gpurun_out/plain_$2.log 2>&1; }
run "C2 kmeans k=64 F=128" c2 && run "C3 multi" c3 && run "C5 MLP" mlp
N="ncu --set full --clock-control none --import-source on -f"
$N -k regex:regression_tile_kernel -s 3 -c 1 -o gpurun_out/ncu_c2_r02 python tools/bench_configs.py --only "C2 kmeans k=64 F=128" > gpurun_out/ncu_c2.log 2>&1
$N -k regex:regression_tile_kernel -s 6 -c 1 -o gpurun_out/ncu_c3_r02 python tools/bench_configs.py --only "C3 multi" > gpurun_out/ncu_c3.log 2>&1
$N -k regex:sg_layer_kernel -s 15 -c 3 -o gpurun_out/ncu_mlp_layer_r02 python tools/bench_configs.py --only "C5 MLP" > gpurun_out/ncu_mlp1.log 2>&1
$N -k regex:sg_slice_inputs_kernel -s 5 -c 1 -o gpurun_out/ncu_mlp_slice_r02 python tools/bench_configs.py --only "C5 MLP" > gpurun_out/ncu_mlp2.log 2>&1
FJP_SLICED=0 $N -k regex:neural_pair_kernel -s 3 -c 1 -o gpurun_out/ncu_mlp_fp64_r02 python tools/bench_configs.py --only "C5 MLP" > gpurun_out/ncu_mlp3.log 2>&1
# the reports of the kernels.cu kernels carry the whole source (40 MB each): summarise here, keep the text
for k in c2 c3 mlp_layer mlp_slice mlp_fp64; do python tools/ncu_summary.py gpurun_out/ncu_${k}_r02.ncu-rep "tools/ncu_configs.sh" > gpurun_out/ncu_${k}_r02.txt; done
rm -f gpurun_out/ncu_c2_r02.ncu-rep gpurun_out/ncu_c3_r02.ncu-rep gpurun_out/ncu_mlp_fp64_r02.ncu-rep
ls -la gpurun_out/ncu_*_r02.*
