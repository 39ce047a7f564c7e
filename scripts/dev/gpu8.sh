mkdir -p gpurun_out
nvidia-smi -L | head -8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
true
true
timeout 600 $TR bench.py --gpus 8 --no-cpu-baseline > gpurun_out/r1_v7_bench_8gpu.json 2> gpurun_out/r1_v7_bench_8gpu.err
for f in gpurun_out/r1_v7_bench_8gpu*.json; do python -c "
import json,sys; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['value'], d['n_gpus'], d['e2e']['value'], d['ms_per_step'])"; done
