#!/bin/bash
# Timing experiments for the assign kernel (results are invalid with knockout != 0): which role limits it?
ROWS=${ROWS:-20000000}
OPTS=${OPTS:-"knockout=0 knockout=4 knockout=64 knockout=2 knockout=128"}   # 4 no epilogue work, 64 TMEM loads only, 2 no transform stores, 128 stores of zeros
for opt in $OPTS; do
  echo "== $opt"
  timeout -k 5 120 python bench.py --rows $ROWS --steps 3 --warmup 3 --no-e2e --no-cpu --option $opt --option role_prof=1 2>&1 | grep -E "role_prof|^\{" | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line)
        print({k: round(v, 2) for k, v in d['phases_ms_per_step'].items()}, 'rechecked', d['rechecked_rows_per_step'], 'exec TF/s', round(3 * d['roofline']['achieved'], 1))
    else:
        print(line.strip())
"
done
