# one-off tuning helper: cost factors of the decoder CTA split (nsb_decoder.cu: NSB_XSAVED_COST, NSB_XFWD_COST,
# NSB_SAVED_COST; NSB_BWD_SHARES / NSB_FWD_SHARES force a split)
run() { env "$@" python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-tracking 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$*', 'step', round(d['ms_per_step'],4), 'bwd', round(d['roofline']['kernel_ms'],4), 'fwd', round(d['roofline']['forward_kernel_ms'],4), 'dw', round(d['roofline']['dweight_kernel_ms'],4))"; }
for c in ${XS:-0.5 0.55 0.6 0.65}; do run NSB_XSAVED_COST=$c NSB_XFWD_COST=1.5; done
for c in ${XF:-1.7 2.0 2.3 2.7}; do run NSB_XSAVED_COST=0.6 NSB_XFWD_COST=$c; done
