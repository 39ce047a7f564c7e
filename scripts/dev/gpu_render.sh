timeout 600 python -m pytest tests/test_gpu_render.py -x -q 2>&1 | tail -3
timeout 200 python scripts/bench_render.py; timeout 200 python scripts/bench_render.py --cameras 2
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_render -c 12 --csv python scripts/bench_render.py --reps 3 2>&1 | grep -i "k_render" | tail -6
