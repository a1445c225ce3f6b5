for rep in 1 2; do
for pr in high same low; do
  ZL_B200_PLAN_PRIORITY=$pr timeout 200 python bench.py --steps 10 --warmup 3 --no-latency --no-cpu-baseline --no-thumb > gpurun_out/s.log 2>&1
  python -c "
import json
d=json.loads(open('gpurun_out/s.log').read().strip().splitlines()[-1]);print('prio $pr', round(d['ms_per_step'],2),round(d['roofline']['kernel_ms'],3),round(d['roofline']['frac'],4))"
done
done
