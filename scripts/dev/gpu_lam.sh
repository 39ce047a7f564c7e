set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for cfg in "12 1" "8 1" "16 1" "6 1" "24 1" "12 0"; do set -- $cfg
BCP_HANDOFF=$1 BCP_SIDE=$2 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_lam.json 2> gpurun_out/bench_lam.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_lam.json').read().strip().splitlines()[-1]); print('LAM handoff $1 side $2', d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d['ms_per_step'])"
done
BCP_LAMBDA=0 timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench_nolam.json 2> gpurun_out/bench_nolam.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_nolam.json').read().strip().splitlines()[-1]); print('OLD', d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d['ms_per_step'])"
