#!/bin/bash
# r02 batch 15 (gpurun --gpus 2): span-mix GPU test, weak / span-mix / strong scaling bench lines at N = 2.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 900 python -m pytest tests/test_spanmix_gpu.py -m gpu -q > $O/b15_spanmix_test.log 2>&1; echo "spanmix test rc=$?" > $O/b15_summary.txt
tail -4 $O/b15_spanmix_test.log >> $O/b15_summary.txt
timeout 900 $T bench.py --gpus 2 > $O/b15_n2.json 2> $O/b15_n2.err; echo "n2 rc=$?" >> $O/b15_summary.txt
timeout 900 $T bench.py --gpus 2 --span-mix 2 > $O/b15_n2_span2.json 2> $O/b15_n2_span2.err; echo "n2 span rc=$?" >> $O/b15_summary.txt
timeout 900 $T bench.py --gpus 2 --sessions 2048 > $O/b15_n2_strong.json 2> $O/b15_n2_strong.err; echo "n2 strong rc=$?" >> $O/b15_summary.txt
python - <<'PY' >> $O/b15_summary.txt
import json
for t in ("n2", "n2_span2", "n2_strong"):
    try:
        j=json.loads(open(f"gpurun_out/b15_{t}.json").read().strip().splitlines()[-1])
        print(f'{t:10s} ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} e2e_pcm16 {j["e2e_pcm16"]["value"]:.4g}')
    except Exception as ex: print(t, "parse failed", ex)
PY
nvidia-smi topo -m > $O/b15_topo.txt 2>&1
cat $O/b15_summary.txt
