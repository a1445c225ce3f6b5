python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for rep in 1 2 3; do
  timeout 200 python bench.py --steps 10 --warmup 3 --no-latency --no-cpu-baseline --no-thumb > gpurun_out/s.log 2>&1
  python -c "
import json
d=json.loads(open('gpurun_out/s.log').read().strip().splitlines()[-1]);print(round(d['ms_per_step'],2),round(d['roofline']['kernel_ms'],3),f\"{d['value']:.4g} {d['e2e']['value']:.4g}\", d['gpu_launches'])"
done
