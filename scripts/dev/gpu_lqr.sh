mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_lqr.py -x -q 2>&1 | tail -25
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_nvml.json 2> gpurun_out/bench_nvml.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_nvml.json').read().strip().splitlines()[-1]); print('B', d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d['ms_per_step'], d['clocks'])"
tail -3 gpurun_out/bench_nvml.err
