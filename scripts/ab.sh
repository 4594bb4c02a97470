#!/bin/bash
# A/B of the pull-at-emit switch at two depths (throughput in M examples/s, ms per step)
for e in 0 1; do for d in 3 4; do
  echo -n "emit $e depth $d: "
  WH_PULL_AT_EMIT=$e python bench.py --steps 400 --depth $d 2>&1 | tail -1 | python -c "import json,sys; l=json.loads(sys.stdin.read()); print(round(l['value']/1e6,1), l['ms_per_step'])"
done; done
