#!/bin/bash
# r02 batch 48: plan_kernel is dispatched ahead of the gather it runs "beside" (trace: its CTAs occupy the 89 us BEFORE the first
# gather CTA starts), so its stand-alone time counts 1:1.  Register cap of plan_kernel (ZL_PLAN_MINB 6/8/10/12/16 = 80/64/48/40/32
# registers): stand-alone duration under ncu and the step.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b48_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
for m in 6 8 10 12 16; do
  L=$PWD/build/v_pm$m.so; [ $m = 6 ] && L=$PWD/libzl_b200/libzl_b200.so
  ZL_B200_LIB=$L timeout 600 python bench.py $S > $O/b48_$m.json 2> $O/b48_$m.err
  ZL_B200_LIB=$L timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"plan_kernel" -s 40 -c 4 --csv --log-file $O/b48_plan_$m.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b48_ncu_$m.log 2>&1
  python - $m <<'PY' >> $O/b48_summary.txt
import json,sys,csv
m=sys.argv[1]
j=json.loads(open(f"gpurun_out/b48_{m}.json").read().strip().splitlines()[-1]); r=j["roofline"]; p=j.get("parity_at_size") or {}
d=[float(row[-1].replace(",","")) for row in csv.reader(l for l in open(f"gpurun_out/b48_plan_{m}.csv") if l.startswith('"')) if "plan_kernel" in ",".join(row)]
print(f'MINB {m}: ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} parity {p.get("masters_bit_equal")} plan_kernel us {[round(x/1000,1) for x in d]}')
PY
done
cat $O/b48_summary.txt
