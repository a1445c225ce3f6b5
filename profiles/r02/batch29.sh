#!/bin/bash
# r02 batch 29: event timeline of the shipped schedule; live gather without the overlapped port mix.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O; : > $O/b29_summary.txt
S="--steps 6 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity"
rm -f $O/b29_timeline.txt; ZL_B200_TIMELINE=$O/b29_timeline.txt timeout 300 python bench.py $S --no-kernel-events > $O/b29_tl.json 2> $O/b29_tl.err; echo "timeline rc=$?" >> $O/b29_summary.txt
ZL_B200_MIX_OVERLAP=0 timeout 300 python bench.py --steps 12 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b29_nomixovl.json 2> $O/b29_nomixovl.err; echo "nomixovl rc=$?" >> $O/b29_summary.txt
python - <<'PY' >> $O/b29_summary.txt
import json
for n in ("nomixovl",):
    try:
        j=json.loads(open(f"gpurun_out/b29_{n}.json").read().strip().splitlines()[-1]); r=j["roofline"]
        print(f'{n}: ms/step {j["ms_per_step"]:.3f} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f}')
    except Exception as ex: print(n, "parse failed", ex)
PY
cat $O/b29_summary.txt
