#!/bin/bash
# ncu evidence for the hot-path rows b/c/d (run on the GPU box via gpurun; outputs in gpurun_out/).
# 1) launch list of our kernels in one bench_paths.py --quick run; 2) one --set full capture per kernel.
set -u
OUT=gpurun_out
TAG=${1:-r01}
mkdir -p $OUT
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -c 3000 --csv \
    --log-file $OUT/${TAG}_launches_paths.csv python bench_paths.py --quick > $OUT/ncu_paths_list.log 2>&1
for K in k_gemm3x_nt k_gemv_rows k_sqnorm_proj k_hinge_partial k_pr_rda k_sqnorm_tc k_syrk_f16; do
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:^${K} --launch-skip 3 -c 2 -f \
        -o $OUT/${TAG}_${K} python bench_paths.py --quick > $OUT/ncu_${K}.log 2>&1
    if [ -f $OUT/${TAG}_${K}.ncu-rep ]; then
        ncu -i $OUT/${TAG}_${K}.ncu-rep --page raw --csv > $OUT/${TAG}_${K}_raw.csv 2>/dev/null
    fi
done
ls -la $OUT
