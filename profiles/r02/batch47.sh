#!/bin/bash
# r02 batch 47: what would fewer hand-over bytes buy?  Experiment build (results WRONG by design): fast-path voices write only
# tile 0's descriptor, tiles 1-3 keep whatever the buffer held (stale descriptors of the same voice two periods earlier), so the
# gather's work stays statistically the same while plan_kernel writes 100 MB less per period.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b47_summary.txt
S="--steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
for v in 0 1; do
  ZL_B200_EXP_FEWDESCS=$v ZL_B200_LIB=$PWD/build/v_exp.so timeout 600 python bench.py $S > $O/b47_$v.json 2> $O/b47_$v.err
  python - $v <<'PY' >> $O/b47_summary.txt
import json,sys
j=json.loads(open(f"gpurun_out/b47_{sys.argv[1]}.json").read().strip().splitlines()[-1]); r=j["roofline"]
print(f'fewdescs {sys.argv[1]}: ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} vf/step {j["voice_frames_per_step"]:.4g}')
PY
done
ZL_B200_EXP_FEWDESCS=1 ZL_B200_LIB=$PWD/build/v_exp.so timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -k regex:"plan_kernel" -s 40 -c 3 --csv --log-file $O/b47_plan.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b47_ncu.log 2>&1
grep -E "plan_kernel" $O/b47_plan.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | head -12 >> $O/b47_summary.txt
T="--steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity --no-kernel-events"
ZL_B200_EXP_FEWDESCS=1 ZL_B200_LIB=$PWD/build/v_exp_trace.so ZL_B200_CTA_TRACE=$O/b47_trace.txt timeout 300 python bench.py $T > $O/b47_live.json 2> $O/b47_live.err
python profiles/r02/cta_trace.py $O/b47_trace.txt >> $O/b47_summary.txt 2>&1
cat $O/b47_summary.txt
