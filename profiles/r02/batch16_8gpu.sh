#!/bin/bash
# r02 batch 16 (gpurun --gpus 8): weak / span-mix / strong scaling bench lines at N = 8 and N = 4.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
: > $O/b16_summary.txt
run() { tag=$1; n=$2; shift 2; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n "$@" > $O/b16_$tag.json 2> $O/b16_$tag.err; echo "$tag rc=$?" >> $O/b16_summary.txt; }
run n8 8
run n8_span8 8 --span-mix 8
run n8_strong 8 --sessions 512
run n4 4
run n4_span4 4 --span-mix 4
run n4_strong 4 --sessions 1024
python - <<'PY' >> $O/b16_summary.txt
import json
for t in ("n8", "n8_span8", "n8_strong", "n4", "n4_span4", "n4_strong"):
    try:
        j=json.loads(open(f"gpurun_out/b16_{t}.json").read().strip().splitlines()[-1])
        print(f'{t:10s} ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e {j["e2e"]["value"]:.4g} e2e_pcm16 {j["e2e_pcm16"]["value"]:.4g}')
    except Exception as ex: print(t, "parse failed", ex)
PY
nvidia-smi topo -m > $O/b16_topo.txt 2>&1; (lscpu | head -25; free -g) > $O/b16_host.txt 2>&1
cat $O/b16_summary.txt
