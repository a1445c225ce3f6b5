#!/bin/bash
# r02 batch 12: why voices leave the planner's fast path (counters); SyncTimer facade test.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
ZL_B200_WALK_PROF=1 timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity > $O/b12_prof.json 2> $O/b12_prof.err; grep "walk prof" $O/b12_prof.err > $O/b12_summary.txt
timeout 1200 python -m pytest tests/test_facade.py tests/test_sched.py -m gpu -q > $O/b12_pytest.log 2>&1; echo "pytest rc=$?" >> $O/b12_summary.txt
tail -12 $O/b12_pytest.log >> $O/b12_summary.txt
cat $O/b12_summary.txt
