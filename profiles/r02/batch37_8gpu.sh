#!/bin/bash
# r02 batch 37 (gpurun --gpus 8): final build - span-mix GPU test (2 ranks), weak / span-mix / strong scaling bench lines at N = 8, 4, 2.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
: > $O/b37_summary.txt
CUDA_VISIBLE_DEVICES=0,1 timeout 900 python -m pytest tests/test_spanmix_gpu.py -m gpu -q > $O/b37_spanmix_test.log 2>&1; echo "spanmix test rc=$?" >> $O/b37_summary.txt
tail -3 $O/b37_spanmix_test.log >> $O/b37_summary.txt
run() { tag=$1; n=$2; shift 2; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --no-thumb "$@" > $O/b37_$tag.json 2> $O/b37_$tag.err; echo "$tag rc=$?" >> $O/b37_summary.txt; }
run n8 8
run n8_span8 8 --span-mix 8
run n8_strong 8 --sessions 512
run n4 4
run n4_span4 4 --span-mix 4
run n4_strong 4 --sessions 1024
run n2 2
run n2_span2 2 --span-mix 2
run n2_strong 2 --sessions 2048
python - <<'PY' >> $O/b37_summary.txt
import json
for t in ("n8", "n8_span8", "n8_strong", "n4", "n4_span4", "n4_strong", "n2", "n2_span2", "n2_strong"):
    try:
        j=json.loads(open(f"gpurun_out/b37_{t}.json").read().strip().splitlines()[-1])
        print(f'{t:10s} ms/step {j["ms_per_step"]:.3f} vf/s {j["value"]:.4g} e2e(pcm16) {j["e2e"]["value"]:.4g} e2e_f32 {j["e2e_f32"]["value"]:.4g}')
    except Exception as ex: print(t, "parse failed", ex)
PY
cat $O/b37_summary.txt
