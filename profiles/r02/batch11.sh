#!/bin/bash
# r02 batch 11: PCM read-back tests + bench with e2e_pcm16; walk_kernel per-category cycle profile.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_wav.py tests/test_thumbnail_gpu.py -m gpu -q > $O/b11_pytest.log 2>&1; echo "pytest rc=$?" > $O/b11_summary.txt
tail -4 $O/b11_pytest.log >> $O/b11_summary.txt
timeout 900 python bench.py > $O/b11_bench.json 2> $O/b11_bench.err; echo "bench rc=$?" >> $O/b11_summary.txt
python - <<'PY' >> $O/b11_summary.txt
import json
try:
    j=json.loads(open("gpurun_out/b11_bench.json").read().strip().splitlines()[-1]); r=j["roofline"]
    print(f'ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} e2e_pcm16 {j["e2e_pcm16"]["value"]:.4g} gather_live {r["kernel_ms"]:.4f} alone {r["kernel_alone_ms"]:.4f} frac {r["frac"]:.3f}')
    print("wav_io", j.get("wav_io")); print("parity", j.get("parity_at_size", {}).get("masters_bit_equal"))
except Exception as ex: print("bench parse failed", ex)
PY
ZL_B200_WALK_PROF=1 timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b11_prof.json 2> $O/b11_prof.err; grep "walk prof" $O/b11_prof.err >> $O/b11_summary.txt
cat $O/b11_summary.txt
