#!/bin/bash
# r02 batch 1 (GPU box): full GPU suite on the default build, bench, kernel variants, planner-cost decomposition.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/b1_gpu.txt
B="python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-latency --no-thumb"
sum() { python - "$1" "$2" <<'PY'
import json,sys
try:
    j=json.loads(open(sys.argv[2]).read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'{sys.argv[1]:28s} ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f}')
except Exception as ex:
    print(sys.argv[1], "FAILED", ex)
PY
}
timeout 1500 python -m pytest tests -m gpu -x -q > $O/b1_pytest.log 2>&1; echo "pytest rc=$?" > $O/b1_summary.txt
tail -3 $O/b1_pytest.log >> $O/b1_summary.txt
timeout 300 $B > $O/b1_default.json 2> $O/b1_default.err; sum default $O/b1_default.json >> $O/b1_summary.txt
for v in ldgsts_s2m28 tma_s3m28 tma_s2m24 tma_s2m32 kt256_s2m20 kt256_s2m24; do
  ZL_B200_LIB=$PWD/build/variants/$v.so timeout 300 $B > $O/b1_$v.json 2> $O/b1_$v.err; sum $v $O/b1_$v.json >> $O/b1_summary.txt
done
ZL_B200_DEBUG_SKIP=3 timeout 300 $B > $O/b1_skip3.json 2> $O/b1_skip3.err; sum "default skip plan+walk" $O/b1_skip3.json >> $O/b1_summary.txt
ZL_B200_DEBUG_SKIP=2 timeout 300 $B > $O/b1_skip2.json 2> $O/b1_skip2.err; sum "default skip walk" $O/b1_skip2.json >> $O/b1_summary.txt
ZL_B200_PLAN_PRIORITY=low timeout 300 $B > $O/b1_prio_low.json 2> $O/b1_prio_low.err; sum "default planner low prio" $O/b1_prio_low.json >> $O/b1_summary.txt
ZL_B200_PLAN_PRIORITY=same timeout 300 $B > $O/b1_prio_same.json 2> $O/b1_prio_same.err; sum "default planner same prio" $O/b1_prio_same.json >> $O/b1_summary.txt
ZL_B200_MIX_OVERLAP=0 timeout 300 $B > $O/b1_nomixovl.json 2> $O/b1_nomixovl.err; sum "default mix not overlapped" $O/b1_nomixovl.json >> $O/b1_summary.txt
# instruction counts + launch list of the default build
timeout 600 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"plan_kernel|walk_kernel|gather_kernel|master_mix" -c 140 --csv --log-file $O/b1_launches.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-latency --no-thumb > $O/b1_ncu.log 2>&1
cat $O/b1_summary.txt
