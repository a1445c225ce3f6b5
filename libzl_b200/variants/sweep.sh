for rep in 1 2 3; do
for cfg in "0 0" "2 4" "3 4" "4 6" "6 8"; do
  set -- $cfg
  ZL_B200_PLAN_CTAS=$1 ZL_B200_WALK_CTAS=$2 timeout 200 python bench.py --steps 10 --warmup 3 --no-latency --no-cpu-baseline > gpurun_out/s.log 2>&1
  python -c "
import json
d=json.loads(open('gpurun_out/s.log').read().strip().splitlines()[-1]);print('plan $1 walk $2', round(d['ms_per_step'],2),round(d['roofline']['kernel_ms'],3),round(d['roofline']['frac'],4))"
done
done
