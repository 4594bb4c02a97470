#!/bin/bash
# One-GPU timings of BASELINE.json's other configurations (parity-test cases, not bench lines).
for a in "--algo adagrad --loss logit" "--algo sgd --loss sqhinge" "--shape kdd --pool 4 --steps 60"; do
  echo "== $a"; python bench.py --steps 200 --warmup 10 $a 2>&1 | tail -1 | python -c "
import json,sys
l=json.loads(sys.stdin.read()); print(json.dumps({k:l[k] for k in ('value','ms_per_step','e2e')}), l['config']['unique_keys_per_row_r'], l['step_roofline'], l['kernel_ms'])"
done
