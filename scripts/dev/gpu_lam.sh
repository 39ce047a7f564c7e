mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for i in 1 2; do
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lam.json 2> gpurun_out/bench_lam.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_lam.json').read().strip().splitlines()[-1]); print('LAM', d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d['ms_per_step'])"
done
