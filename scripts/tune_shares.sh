# one-off tuning helper: CTA split of the 3-decoder stage (NSB_BWD_SHARES / NSB_FWD_SHARES hooks in nsb_decoder.cu)
for sh in ${SHARES:-"22,25,101" "21,24,103" "20,24,104" "21,23,104" "20,22,106" "23,25,100" "22,24,102"}; do
  NSB_BWD_SHARES=$sh python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-tracking 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('bwd $sh', round(d['ms_per_step'],4), round(d['roofline']['kernel_ms'],4), round(d['roofline']['forward_kernel_ms'],4), round(d['roofline']['dweight_kernel_ms'],4))"
done
