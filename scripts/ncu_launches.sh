# usage: bash scripts/ncu_launches.sh <out.csv> [bench args...]   (after the same command exited 0 without ncu)
out=$1; shift
python bench.py "$@" > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file $out python bench.py "$@" > gpurun_out/ncu.log 2>&1
