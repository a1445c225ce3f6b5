#!/bin/bash
# r02 batch 14: zl_levels_scan_static alone in a fresh process, then the whole GPU suite (it is collected last), then the bench.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/b14_noalloc_alone.log 2>&1; echo "no_alloc alone rc=$?" > $O/b14_summary.txt
tail -3 $O/b14_noalloc_alone.log >> $O/b14_summary.txt
timeout 2400 python -m pytest tests -m gpu -q -x > $O/b14_pytest.log 2>&1; echo "pytest rc=$?" >> $O/b14_summary.txt
tail -6 $O/b14_pytest.log >> $O/b14_summary.txt
timeout 900 python bench.py > $O/b14_bench.json 2> $O/b14_bench.err; echo "bench rc=$?" >> $O/b14_summary.txt
python - <<'PY' >> $O/b14_summary.txt
import json
try:
    j=json.loads(open("gpurun_out/b14_bench.json").read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} e2e_pcm16 {j["e2e_pcm16"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f} whole {r["whole_step_frac"]:.3f}')
    print("latency", j.get("latency")); print("parity", (j.get("parity_at_size") or {}).get("masters_bit_equal"))
except Exception as ex: print("bench parse failed", ex)
PY
tail -3 $O/b14_bench.err >> $O/b14_summary.txt
cat $O/b14_summary.txt
