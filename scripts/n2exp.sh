run() { echo "== $*"; env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 60 --warmup 10 --no-parity --no-others --no-cpu $EXTRA 2>/dev/null | python -c "import sys,json; j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(j['value']/1e9,4),'G ex/s', round(j['ms_per_step'],4),'ms', j['kernel_ms'])"; }
run WH_NOP=1
run WH_APPLY_BLOCKS=3
run WH_APPLY_BLOCKS=4
run WH_FUSED_SPARE_SMS=8
