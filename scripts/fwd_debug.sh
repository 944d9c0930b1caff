for d in 0 1 2 3; do
  NSB_DEBUG_FWD=$d timeout 300 python bench.py --quick --no-cpu-baseline 2>/dev/null | grep "^{" | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['roofline']; print('debug $d', 'fwd', round(r['kernel_ms']*1e3,1), 'step', round(d['ms_per_step']*1e3,1))"
done
