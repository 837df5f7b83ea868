#!/bin/bash
# ncu evidence for the sign-prox sweep iteration and the tensor-path comp-uprjdists builder
# (run on the GPU box via gpurun; outputs in gpurun_out/).
set -u
OUT=gpurun_out
TAG=${1:-r01}
mkdir -p $OUT
timeout 300 python tools/prof_new.py > $OUT/prof_new_plain.log 2>&1 || { echo "prof_new failed"; tail -5 $OUT/prof_new_plain.log; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_' --csv \
    --log-file $OUT/${TAG}_launches_new.csv python tools/prof_new.py > $OUT/ncu_new_list.log 2>&1
# third sweep iteration + the second builder pass: skip the warm-up launches of each kernel family
timeout 1500 ncu --set full --clock-control none --import-source on \
    -k regex:'k_gemm3x_nt|k_sign_prep|k_quad_coldot|k_pool_fulldist|k_uprj_finish|k_desc' -s 70 -c 70 -f \
    -o $OUT/${TAG}_new python tools/prof_new.py > $OUT/ncu_new_full.log 2>&1
ncu -i $OUT/${TAG}_new.ncu-rep --page raw --csv > $OUT/${TAG}_new_raw.csv 2>/dev/null
rm -f $OUT/${TAG}_new.ncu-rep
ls -la $OUT | tail -8
