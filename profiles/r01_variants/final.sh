python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; tail -c 300 gpurun_out/final_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference.json 2>&1; tail -c 200 gpurun_out/final_bench_reference.json
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"plan_kernel|walk_kernel|gather_kernel|master_mix|thumb|wav_|interleave|jatm" --csv --log-file gpurun_out/final_launches.csv python bench.py --steps 2 --warmup 3 --no-latency --no-cpu-baseline > gpurun_out/final_ncu.log 2>&1; tail -c 200 gpurun_out/final_ncu.log
