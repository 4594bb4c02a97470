# usage: bash scripts/ncu_full.sh <out-basename> <kernel-regex> <skip> <count> [bench args...]
out=$1; re=$2; skip=$3; cnt=$4; shift 4
python bench.py "$@" > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"$re" -s $skip -c $cnt -o gpurun_out/$out python bench.py "$@" > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log | cut -c1-200
