mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
run() { timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline $2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['value'], d['kernel_ms_per_step'])"; }
for c in 50 6 8 10 12 16; do BCP_SOLVE_CAP=$c run "cap $c" ""; done
BCP_SOLVE_CAP=8 BCP_FETCH_MIN=4 run "cap 8 fm 4" ""
BCP_SOLVE_CAP=8 BCP_FETCH_MIN=16 run "cap 8 fm 16" ""
