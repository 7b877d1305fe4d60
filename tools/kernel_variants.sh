#!/bin/bash
# assign-kernel variants timed on the same workload (C3 shape, ROWS rows)
ROWS=${ROWS:-20000000}
VARIANTS=${VARIANTS:-"kernel=1|kernel=2 cluster=1|kernel=2 cluster=3|kernel=2 cluster=2"}
IFS='|' read -ra VS <<< "$VARIANTS"
for v in "${VS[@]}"; do
  echo "== $v"
  opts=""; for o in $v; do opts="$opts --option $o"; done
  timeout -k 5 150 python bench.py --rows $ROWS --steps 3 --warmup 3 --no-e2e --no-cpu $opts --option role_prof=1 2>&1 | grep -E "role_prof|^\{|rror" | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line)
        print({k: round(v, 2) for k, v in d['phases_ms_per_step'].items()}, 'rechecked', d['rechecked_rows_per_step'], 'full', d.get('full_rechecks_per_step'), 'exec TF/s', round(3 * d['roofline']['achieved'], 1), 'cost', d['final_cost'])
    else:
        print(line.strip())
"
done
