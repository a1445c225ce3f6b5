# usage: bash libzl_b200/variants/runv.sh a b c ...   (on the GPU box; each variant = one .so built with different -D)
for rep in 1 2; do
for v in "$@"; do
  ZL_B200_LIB=$PWD/libzl_b200/variants/v_$v.so timeout 200 python bench.py --steps 10 --warmup 3 --no-latency --no-cpu-baseline --no-thumb > gpurun_out/v_$v.log 2>&1
  python -c "
import json,sys
try:
  d=json.loads(open('gpurun_out/v_$v.log').read().strip().splitlines()[-1]);print('$v',round(d['ms_per_step'],2),round(d['roofline']['kernel_ms'],3),round(d['roofline']['kernel_alone_ms'],3))
except Exception as e: print('$v','fail',e)"
done
done
