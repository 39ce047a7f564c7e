run() { timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline $2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['value'], d['kernel_ms_per_step'], d['e2e']['value'])"; }
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run "65536" ""
run "262144" "--envs 262144"
