O=gpurun_out
for n in 8 4 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 20 --warmup 3 > $O/r2_bench_${n}gpu.json 2> $O/r2_bench_${n}gpu.err; echo "bench $n rc=$?"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --impl reference --gpus 8 --steps 5 --warmup 1 > $O/r2_bench_reference_8gpu.json 2> $O/r2_bench_reference_8gpu.err; echo "ref rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29532 scripts/accuracy_throughput_sweep.py --envs 1048576 --steps 20 > $O/r2_accuracy_throughput_8gpu.json 2> $O/r2_accuracy_throughput_8gpu.err; echo "sweep rc=$?"
nproc; free -g | head -2
