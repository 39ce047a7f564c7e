mkdir -p gpurun_out
for envs in 1048576 262144 131072; do
for lam in 2 0; do
BCP_LAMBDA=$lam timeout 600 python bench.py --envs $envs --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_x.json').read().strip().splitlines()[-1]); print('envs $envs lam $lam', d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d['ms_per_step'])"
done; done
BCP_LAMBDA=2 timeout 900 python -m pytest tests/test_gpu_split.py tests/test_gpu_parity.py tests/test_gpu_protocol.py -m gpu -x -q 2>&1 | tail -3
