#!/bin/bash
# Round-2 evidence for the headline path (run on the GPU box via gpurun; outputs in gpurun_out/):
#   1) the bench line itself (no profiler), 2) the ncu launch list of a short run of the same command
#   (gpu__time_duration per launch: the kernels' SHARES of a step), 3) one --set full capture of the dominant kernel
#   (tensor pipe, DRAM bytes: `roofline.traffic` of bench.py is read from this export), 4) the new kernels of round 2.
set -u
OUT=gpurun_out
TAG=${1:-r02}
mkdir -p $OUT
timeout 900 python bench.py --steps 3 --warmup 3 > $OUT/${TAG}_bench_500k_n1.json 2> $OUT/${TAG}_bench_500k_n1.err || { tail -5 $OUT/${TAG}_bench_500k_n1.err; exit 1; }
SHORT="python bench.py --steps 1 --warmup 1 --pairs 8192 --no-cpu --no-e2e --no-rows"
timeout 300 $SHORT > $OUT/short_plain.log 2>&1 || { echo "short run failed"; exit 1; }
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/${TAG}_launches_bench_8192pairs.csv $SHORT > $OUT/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_pool_fulldist|k_desc|k_ring_reduce|k_gen_pr_filters' \
    -s 4 -c 4 -f -o $OUT/${TAG}_pool python bench.py --steps 1 --warmup 1 --pairs 8192 --no-cpu --no-e2e --no-rows > $OUT/ncu_full.log 2>&1
ncu -i $OUT/${TAG}_pool.ncu-rep --page raw --csv > $OUT/${TAG}_pool_fulldist_ncu_full.csv 2>/dev/null
rm -f $OUT/${TAG}_pool.ncu-rep
ls -la $OUT | tail -8
