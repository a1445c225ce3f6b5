#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -q > $O/b8_pytest.log 2>&1; echo "pytest rc=$?" > $O/b8_summary.txt
tail -4 $O/b8_pytest.log >> $O/b8_summary.txt
timeout 900 python bench.py > $O/b8_bench.json 2> $O/b8_bench.err; echo "bench rc=$?" >> $O/b8_summary.txt
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/b8_bench_ref.json 2> $O/b8_bench_ref.err; echo "bench ref rc=$?" >> $O/b8_summary.txt
python - <<'PY' >> $O/b8_summary.txt
import json
try:
    j=json.loads(open("gpurun_out/b8_bench.json").read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f}')
    print("parity_at_size", j.get("parity_at_size")); print("latency", j.get("latency"), j.get("cpu_latency")); print("cpu", j.get("cpu_baseline")); print("thumb", j.get("thumbnail")); print("host", j.get("host_submit_ms_per_step"))
except Exception as ex: print("bench parse failed", ex)
PY
timeout 900 compute-sanitizer --tool memcheck --log-file $O/b8_memcheck.log python profiles/r02/sanitize_cases.py > $O/b8_memcheck.out 2>&1; echo "memcheck rc=$?" >> $O/b8_summary.txt; tail -3 $O/b8_memcheck.log >> $O/b8_summary.txt
timeout 1200 compute-sanitizer --tool racecheck --log-file $O/b8_racecheck.log python profiles/r02/sanitize_cases.py > $O/b8_racecheck.out 2>&1; echo "racecheck rc=$?" >> $O/b8_summary.txt; tail -3 $O/b8_racecheck.log >> $O/b8_summary.txt
cat $O/b8_summary.txt
