O=gpurun_out
for n in 8 4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 3 > $O/r2_bench_${n}gpu.json 2> $O/r2_bench_${n}gpu.err; echo "bench $n rc=$?"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 8 --steps 20 --warmup 3 --e2e-parts 1 --no-configs3 > $O/r2_bench_8gpu_parts1.json 2> /dev/null; echo "parts1 rc=$?"
