# usage: gpu_ab.sh VAR a b  -> bench with VAR=a / VAR=b alternating
mkdir -p gpurun_out
V=$1; shift
for x in $1 $2 $1 $2; do
env $V=$x timeout 600 python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/bench_ab.json 2> gpurun_out/bench_ab.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_ab.json').read().strip().splitlines()[-1]); print('$V=$x', d['value'], d['e2e']['value'], d['kernel_ms_per_step'], d['ms_per_step'])"
done
