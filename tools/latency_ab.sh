#!/bin/bash
# Latency-path A/B: cfg1 / cfg2 (one small scenario) and a 1024-agent batch with the 32-lane kernel, for every "name:lib".
for spec in "$@"; do
  IFS=: read name lib <<< "$spec"
  export MPCCBF_LIB=$PWD/$lib; [ "$lib" = "-" ] && unset MPCCBF_LIB
  for cfg in "1 2 16" "1 8 20" "64 16 20"; do
    set -- $cfg
    python bench.py --scenarios $1 --agents $2 --k-hor $3 --steps 200 --warmup 20 --no-extras --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); e = d.get('e2e') or {}
print('$name S=$1 N=$2 K=$3: device %.4g QPs/s  p50 step %.1f us  e2e %.4g' % (d['value'], 1e3 * d.get('p50_step_ms', 0), e.get('value', 0)))"
  done
done
