mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_host_step.py -m gpu -x -q 2>&1 | tail -15
timeout 300 python scripts/dev/e2e_breakdown.py
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_hs.json 2> gpurun_out/bench_hs.err; tail -2 gpurun_out/bench_hs.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_hs.json').read().strip().splitlines()[-1]); print('HS', d['value'], d['e2e'], d['ms_per_step'])"
