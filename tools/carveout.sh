for co in -1 100 90 86 85; do
  MPCCBF_THREAD_CARVEOUT=$co python bench.py --steps 48 --warmup 8 --device-only --no-extras --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print('carveout=$co  %.4f ms  %.4g QPs/s' % (d['kernel_ms'], d['value']))"
done
