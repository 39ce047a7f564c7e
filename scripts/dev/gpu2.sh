mkdir -p gpurun_out
run() { timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline $2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['value'], d['kernel_ms_per_step'])"; }
run base ""
BCP_SOLVE_BLOCKS_PER_SM=1 run "1 block/SM" ""
for fm in 4 16 24; do BCP_FETCH_MIN=$fm run "fetch_min $fm" ""; done
BCP_SOLVE_BLOCKS_PER_SM=1 BCP_FETCH_MIN=4 run "1 block/SM fm4" ""
BCP_SOLVE_BLOCKS_PER_SM=1 run "1 block/SM 262144" "--envs 262144"
run "262144" "--envs 262144"
