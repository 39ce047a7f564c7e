mkdir -p gpurun_out
for envs in 131072 32768 65536; do
for lam in 1 0; do
BCP_HANDOFF=8 BCP_LAMBDA=$lam timeout 600 python bench.py --envs $envs --initial-force 55 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_x.json 2> gpurun_out/bench_x.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_x.json').read().strip().splitlines()[-1]); print('envs $envs lam $lam if55', d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d['ms_per_step'])"
done; done
