#!/bin/bash
# r02 batch 51: confirmation of the final round-2 build (persistent plan_kernel grid as default), smoke(): full GPU suite, bench both arms, launch list, ncu --set full of the gather.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
python -c "import __graft_entry__ as g; g.smoke()" > $O/b51_smoke.log 2>&1; echo "smoke rc=$?" > $O/b51_smoke_rc.txt
timeout 2400 python -m pytest tests -m gpu -q > $O/b51_pytest.log 2>&1; echo "pytest rc=$?" > $O/b51_summary.txt
tail -6 $O/b51_pytest.log >> $O/b51_summary.txt
timeout 900 python bench.py > $O/b51_bench.json 2> $O/b51_bench.err; echo "bench rc=$?" >> $O/b51_summary.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/b51_bench_ref.json 2> $O/b51_bench_ref.err; echo "bench ref rc=$?" >> $O/b51_summary.txt
python - <<'PY' >> $O/b51_summary.txt
import json
try:
    j=json.loads(open("gpurun_out/b51_bench.json").read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f}')
    print("parity_at_size", j.get("parity_at_size")); print("latency", j.get("latency"), j.get("cpu_latency")); print("cpu", j.get("cpu_baseline")); print("host", j.get("host_submit_ms_per_step"))
except Exception as ex: print("bench parse failed", ex)
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/b51_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b51_ncu_list.log 2>&1; echo "ncu list rc=$?" >> $O/b51_summary.txt
timeout 900 ncu --set full --import-source on --clock-control none -k regex:"gather_kernel" -s 40 -c 1 -o $O/b51_gather_full -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b51_ncu_gather.log 2>&1; echo "ncu full rc=$?" >> $O/b51_summary.txt
cat $O/b51_summary.txt
