#!/bin/bash
# r02 batch 53: plan_kernel prefetches its voice record to L2 before the command phase: stand-alone time (ncu) and step, A/B.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b53_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
for v in base pf base2 pf2; do
  L=$PWD/libzl_b200/libzl_b200.so; case $v in pf*) L=$PWD/build/v_pf.so;; esac
  ZL_B200_LIB=$L timeout 600 python bench.py $S > $O/b53_$v.json 2> $O/b53_$v.err
  python - $v <<'PY' >> $O/b53_summary.txt
import json,sys
j=json.loads(open(f"gpurun_out/b53_{sys.argv[1]}.json").read().strip().splitlines()[-1]); r=j["roofline"]; p=j.get("parity_at_size") or {}
print(f'{sys.argv[1]}: ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} parity {p.get("masters_bit_equal")} {p.get("voice_table_bit_equal")}')
PY
done
for v in base pf; do
  L=$PWD/libzl_b200/libzl_b200.so; [ $v = pf ] && L=$PWD/build/v_pf.so
  ZL_B200_LIB=$L timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"plan_kernel" -s 40 -c 6 --csv --log-file $O/b53_plan_$v.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b53_ncu_$v.log 2>&1
  python - $v <<'PY' >> $O/b53_summary.txt
import csv,sys
d=[float(row[-1].replace(",","")) for row in csv.reader(l for l in open(f"gpurun_out/b53_plan_{sys.argv[1]}.csv") if l.startswith('"')) if "plan_kernel" in ",".join(row)]
print(sys.argv[1], "plan_kernel us", [round(x/1000,1) for x in d])
PY
done
cat $O/b53_summary.txt
