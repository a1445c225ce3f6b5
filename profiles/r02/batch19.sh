#!/bin/bash
# r02 batch 19: register footprint of the co-resident planner kernels (build-time caps), now that the one-warp gather is
# register-limited (28 CTAs/SM).  Variants built by build.build(defines=...) into build/variants/.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
B="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
: > $O/b19_summary.txt
for v in default w16 w24 w32 w24p8 w32p8 g32 g32w32p8; do
  if [ $v = default ]; then unset ZL_B200_LIB; else export ZL_B200_LIB=$PWD/build/variants/v_$v.so; fi
  for rep in 1 2; do
  timeout 300 $B > $O/b19_$v.json 2> $O/b19_$v.err
  python - "$v" $O/b19_$v.json <<'PY' >> $O/b19_summary.txt
import json,sys
try:
    j=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'{sys.argv[1]:12s} ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} e2e {j["e2e"]["value"]:.4g} parity {(j.get("parity_at_size") or {}).get("masters_bit_equal")}')
except Exception as ex:
    print(sys.argv[1], "FAILED", ex)
PY
  done
done
cat $O/b19_summary.txt
