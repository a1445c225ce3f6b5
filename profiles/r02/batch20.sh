#!/bin/bash
# r02 batch 20: ncu range replay of ONE whole zl_render (16 periods, all streams together): overlapped vs one stream.
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out; mkdir -p $O
B="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-latency --no-thumb --no-parity --no-kernel-events"
ZL_B200_PROFILE_RENDER=8 timeout 1200 ncu --replay-mode range --set full --clock-control none -o $O/b20_range_overlap -f $B > $O/b20_overlap.log 2>&1; echo "overlap rc=$?" > $O/b20_summary.txt
ZL_B200_PROFILE_RENDER=8 ZL_B200_NO_OVERLAP=1 timeout 1200 ncu --replay-mode range --set full --clock-control none -o $O/b20_range_serial -f $B > $O/b20_serial.log 2>&1; echo "serial rc=$?" >> $O/b20_summary.txt
tail -5 $O/b20_overlap.log >> $O/b20_summary.txt
ls -la $O/b20_* >> $O/b20_summary.txt
cat $O/b20_summary.txt
