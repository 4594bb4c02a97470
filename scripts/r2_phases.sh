#!/bin/bash
tag=$1
run() { name=$1; shift; env "$@" python bench.py --steps 100 --warmup 10 > gpurun_out/${tag}_${name}.json 2> gpurun_out/${tag}_${name}.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_${name}.json").read().strip().splitlines()[-1])
    print("${name}", round(d["value"]/1e6,1), "M ex/s", round(d["ms_per_step"]*1e3,1), "us/step", {k: round(v*1e3,1) for k,v in d["kernel_ms"].items()})
except Exception as e:
    print("${name} failed", e)
PY
}
run p1 WH_DX_DEBUG_PHASES=1
run p2 WH_DX_DEBUG_PHASES=2
run p3 WH_DX_DEBUG_PHASES=3
run p1s0 WH_DX_DEBUG_PHASES=1 WH_DX_SUPER=0
run p2s0 WH_DX_DEBUG_PHASES=2 WH_DX_SUPER=0
run p3s0 WH_DX_DEBUG_PHASES=3 WH_DX_SUPER=0
