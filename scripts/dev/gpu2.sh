mkdir -p gpurun_out
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --impl reference --gpus $N --steps 5 --warmup 1 > gpurun_out/r1_v10_bench_${N}gpu_reference.json 2> gpurun_out/r1_v10_bench_${N}gpu_reference.err
timeout 600 $TR bench.py --gpus $N --no-cpu-baseline > gpurun_out/r1_v10_bench_${N}gpu.json 2> gpurun_out/r1_v10_bench_${N}gpu.err
tail -2 gpurun_out/r1_v10_bench_${N}gpu.err
for f in gpurun_out/r1_v10_bench_${N}gpu.json gpurun_out/r1_v10_bench_${N}gpu_reference.json; do python -c "
import json,sys; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['value'], d['n_gpus'], d['e2e']['value'], d['ms_per_step'], d.get('clocks'))"; done
