#!/bin/bash
# ncu evidence for the hot-path rows b/c/d (run on the GPU box via gpurun; outputs in gpurun_out/).
# 1) launch list (gpu__time_duration) of our kernels in tools/prof_kernels.py; 2) one --set full
# capture of the same program (our kernels only); raw CSV pages exported next to the report.
set -u
OUT=gpurun_out
TAG=${1:-r01}
mkdir -p $OUT
timeout 300 python tools/prof_kernels.py > $OUT/prof_kernels_plain.log 2>&1 || { echo "prof_kernels failed"; tail -5 $OUT/prof_kernels_plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ --csv \
    --log-file $OUT/${TAG}_launches_kernels.csv python tools/prof_kernels.py > $OUT/ncu_list.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'^k_(gemm3x_nt|gemv|sqnorm|hinge_partial|pr_rda|pj_|syrk)' -f \
    -o $OUT/${TAG}_kernels python tools/prof_kernels.py > $OUT/ncu_full.log 2>&1
ncu -i $OUT/${TAG}_kernels.ncu-rep --page raw --csv > $OUT/${TAG}_kernels_raw.csv 2>/dev/null
# gpurun copies at most 64 MiB back: keep the CSV export, drop the binary report
rm -f $OUT/${TAG}_kernels.ncu-rep
ls -la $OUT | tail -20
