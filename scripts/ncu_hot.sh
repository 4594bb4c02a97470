#!/bin/bash
# full-set capture of the hot-key kernels (after the plain run exited 0)
python scripts/one_step.py 3 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"spmv_fused_hot|loc_mark_hot|loc_rank_hot" -c 3 --launch-skip 3 \
    -o gpurun_out/hot_full -f python scripts/one_step.py 3 > gpurun_out/ncu_hot.log 2>&1
tail -3 gpurun_out/ncu_hot.log
