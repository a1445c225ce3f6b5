set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; tail -c 600 gpurun_out/final_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference.json 2>&1; tail -c 400 gpurun_out/final_bench_reference.json
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"plan_kernel|walk_kernel|gather_kernel|master_mix|thumb|wav_|interleave" --csv --log-file gpurun_out/final_launches.csv python bench.py --steps 2 --warmup 3 --no-latency --no-cpu-baseline > gpurun_out/final_ncu.log 2>&1; tail -c 200 gpurun_out/final_ncu.log
ncu --set full --clock-control none --import-source on -k regex:"wav_decode" -c 2 -o gpurun_out/r01_wav_decode -f python bench.py --steps 1 --warmup 3 --no-latency --no-cpu-baseline > gpurun_out/final_ncu2.log 2>&1; tail -c 200 gpurun_out/final_ncu2.log
ncu --set full --clock-control none --import-source on -k regex:"wav_encode" -c 2 -o gpurun_out/r01_wav_encode -f python bench.py --steps 1 --warmup 3 --no-latency --no-cpu-baseline > gpurun_out/final_ncu3.log 2>&1; tail -c 200 gpurun_out/final_ncu3.log
