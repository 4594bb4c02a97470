#!/bin/bash
# full-set capture of the hot-key kernels (after the plain run exited 0)
python scripts/one_step.py 3 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"${1:-spmv_fused_hot}" -c ${2:-1} --launch-skip ${3:-1} \
    -o gpurun_out/hot_full -f python scripts/one_step.py 3 > gpurun_out/ncu_hot.log 2>&1
tail -3 gpurun_out/ncu_hot.log
