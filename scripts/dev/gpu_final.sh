O=gpurun_out; TAG=r1_v10; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; tail -3 $O/${TAG}_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; tail -1 $O/${TAG}_smoke.log
timeout 300 python scripts/bench_render.py > $O/${TAG}_render_bench.json 2> $O/${TAG}_render_bench.err
timeout 300 python scripts/bench_render.py --cameras 2 >> $O/${TAG}_render_bench.json 2>> $O/${TAG}_render_bench.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render$ -s 2 -c 1 -f -o $O/${TAG}_k_render python scripts/bench_render.py --reps 3 > $O/${TAG}_ncu_render.log 2>&1
timeout 900 python bench.py > $O/${TAG}_bench_f32.json 2> $O/${TAG}_bench_f32.err; tail -c 600 $O/${TAG}_bench_f32.json
