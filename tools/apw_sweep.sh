for S in 4096 2048 1024 512 128; do for apw in 0 32; do
  MPCCBF_THREAD_APW=$apw python bench.py --scenarios $S --steps 48 --warmup 8 --device-only --no-extras --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readline()); print('S=$S apw=$apw value %.4g ms %.4f' % (d['value'], d['ms_per_step']))"
done; done
