#!/bin/bash
# Which fused-kernel variant for which batch: 16-agent scenarios, closed-loop snapshots of bench.py, device leg.
for S in 64 128 256 512 1024; do for lanes in 0 8 32; do
  MPCCBF_LANES=$lanes python bench.py --scenarios $S --steps 48 --warmup 8 --device-only --no-extras --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print('S=$S (%d agents) lanes=$lanes  %.4f ms  %.4g QPs/s' % ($S * 16, d['ms_per_step'], d['value']))"
done; done
