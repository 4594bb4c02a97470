#!/bin/bash
# bench variants: bash scripts/r2_variants.sh <tag>
tag=$1
mkdir -p gpurun_out
run() { name=$1; shift; env "$@" python bench.py --steps 200 --warmup 10 --no-others --no-cpu --no-parity > gpurun_out/${tag}_${name}.json 2> gpurun_out/${tag}_${name}.err; python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/${tag}_${name}.json").read().strip().splitlines()[-1])
    print("${name}", round(d["value"]/1e6,1), "M ex/s", round(d["ms_per_step"]*1e3,1), "us/step", "e2e", round(d["e2e"]["value"]/1e6,1), {k: round(v*1e3,1) for k,v in d["kernel_ms"].items()})
except Exception as e:
    print("${name} failed", e)
PY
}
run staged X=1
run fused WH_DX_STAGED=0

