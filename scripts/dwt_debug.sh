for d in ${DWT_LIST:-0 1 2 4 8 3 7 15}; do
  NSB_DWT_DEBUG=$d timeout 300 python bench.py --quick --only headline --no-cpu-baseline 2>/dev/null | grep "^{" | python -c "
import sys,json
d=json.loads(sys.stdin.read()); r=d['roofline']; print('debug $d', 'dweight', round(r['hbm']['k_dweight_tc']['kernel_ms']*1e3,1), 'step', round(d['ms_per_step']*1e3,1))"
done
