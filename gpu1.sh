set -x
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 3 -c 1 -f -o gpurun_out/prof_r1_v3 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/ncu3.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; tail -3 gpurun_out/bench_a.err; cut -c1-900 gpurun_out/bench_a.json
