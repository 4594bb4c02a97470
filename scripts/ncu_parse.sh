#!/bin/bash
# launch list of the reader kernels (run only after the plain command exited 0)
python scripts/parse_bench.py > gpurun_out/parse_bench.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/parse_launches.csv python scripts/parse_bench.py > gpurun_out/parse_ncu.log 2>&1
