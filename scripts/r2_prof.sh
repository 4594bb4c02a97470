#!/bin/bash
# usage: bash scripts/r2_prof.sh <tag> <kernel-regex> [bench args]: plain run, launch list, one full-set capture of the step kernels
tag=$1; re=$2; shift 2
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 --pool 4 "$@" > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || exit 1
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/${tag}_launches.csv python bench.py --steps 6 --warmup 3 --pool 4 --depth 1 "$@" > gpurun_out/${tag}_ncu.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"$re" -s 20 -c 8 -o gpurun_out/${tag}_full \
    python bench.py --steps 6 --warmup 3 --pool 4 --depth 1 "$@" > gpurun_out/${tag}_ncu_full.log 2>&1
tail -2 gpurun_out/${tag}_ncu_full.log | cut -c1-200
