#!/bin/bash
# r02 batch 22: run-based envelope walk: full GPU suite, bench, walk profile, timeline.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -q -x > $O/b22_pytest.log 2>&1; echo "pytest rc=$?" > $O/b22_summary.txt
tail -6 $O/b22_pytest.log >> $O/b22_summary.txt
timeout 900 python bench.py > $O/b22_bench.json 2> $O/b22_bench.err; echo "bench rc=$?" >> $O/b22_summary.txt
python - <<'PY' >> $O/b22_summary.txt
import json
try:
    j=json.loads(open("gpurun_out/b22_bench.json").read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} e2e_f32 {j["e2e_f32"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f}')
    print("latency", j.get("latency")); print("parity", j.get("parity_at_size"))
except Exception as ex: print("bench parse failed", ex)
PY
tail -3 $O/b22_bench.err >> $O/b22_summary.txt
ZL_B200_WALK_PROF=1 timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b22_prof.json 2> $O/b22_prof.err; grep "walk prof" $O/b22_prof.err >> $O/b22_summary.txt
ZL_B200_NO_OVERLAP=1 timeout 300 python bench.py --steps 8 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b22_serial.json 2> $O/b22_serial.err
python - <<'PY' >> $O/b22_summary.txt
import json
try:
    j=json.loads(open("gpurun_out/b22_serial.json").read().strip().splitlines()[-1]); print("one stream: ms/step", round(j["ms_per_step"], 3))
except Exception as ex: print("serial parse failed", ex)
PY
rm -f $O/b22_timeline.txt; ZL_B200_TIMELINE=$O/b22_timeline.txt timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-kernel-events --no-parity > $O/b22_tl.json 2> $O/b22_tl.err
cat $O/b22_summary.txt
