python -m pytest tests/test_wav.py -x -q 2>&1 | tail -3
python bench.py --steps 2 --warmup 3 --no-latency --no-cpu-baseline > gpurun_out/s.log 2>&1; python -c "
import json
d=json.loads(open('gpurun_out/s.log').read().strip().splitlines()[-1]); print(d['wav_io'])"
ncu --set full --clock-control none --import-source on -k regex:"wav_decode|wav_encode" -c 8 -o gpurun_out/r01_wav -f python bench.py --steps 1 --warmup 3 --no-latency --no-cpu-baseline > gpurun_out/final_ncu2.log 2>&1; tail -c 200 gpurun_out/final_ncu2.log
