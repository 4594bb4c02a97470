#!/bin/bash
# Round-1 profile D: launch list (times + DRAM bytes) and one full-set capture of the chain kernels, on the command
# `python bench.py --steps 3 --warmup 3 --pool 4 --depth 1` (run plainly first; ncu only after it exited 0).
set -e
ARGS="--steps 3 --warmup 3 --pool 4 --depth 1"
python bench.py $ARGS > gpurun_out/d_plain.log 2>&1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv \
    --log-file gpurun_out/d_launches.csv python bench.py $ARGS > gpurun_out/d_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on \
    -k regex:"loc_mark_hot|scan_coop|server_pull|spmv_fused|loss_reduce|server_push|auc_cluster" \
    --launch-skip 35 -c 7 -o gpurun_out/d_full -f python bench.py $ARGS > gpurun_out/d_ncu2.log 2>&1
tail -2 gpurun_out/d_ncu2.log
