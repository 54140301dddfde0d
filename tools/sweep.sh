#!/bin/bash
# usage: tools/sweep.sh "VAR1=a VAR2=b" "VAR1=c" ...   -> one bench line per configuration
for cfg in "$@"; do
  out=$(env $cfg python bench.py --steps 30 --warmup 3 --no-cpu-baseline 2>&1 | tail -1)
  echo "$cfg => $(echo "$out" | python -c 'import json,sys
try:
    d=json.loads(sys.stdin.read()); print("value=%.0f ms=%.3f e2e=%.0f unstable=%d" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["unstable_env_steps"]))
except Exception as e: print("FAILED", e)')"
done
