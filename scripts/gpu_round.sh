#!/bin/bash
# One GPU-box pass that produces every artifact summarised under profiles/:
# tests, smoke, the bench lines (both arms, fp64, upstream-style switches, large batch),
# the ncu launch list of the bench command and one --set full capture per split kernel.
# Usage (from the repo root, on the GPU box):  bash scripts/gpu_round.sh [tag]
TAG=${1:-r1}
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; tail -3 $O/${TAG}_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; tail -3 $O/${TAG}_smoke.log
timeout 900 python bench.py > $O/${TAG}_bench_f32.json 2> $O/${TAG}_bench_f32.err
timeout 900 python bench.py --impl reference > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err
timeout 900 python bench.py --dtype fp64 --no-cpu-baseline > $O/${TAG}_bench_f64.json 2> $O/${TAG}_bench_f64.err
timeout 900 python bench.py --cart-friction 1.0 --steps-per-repeat 5 --initial-force 55 --no-cpu-baseline > $O/${TAG}_bench_upstream_style.json 2> $O/${TAG}_bench_upstream_style.err
timeout 900 python bench.py --envs 262144 --no-cpu-baseline > $O/${TAG}_bench_262144.json 2> $O/${TAG}_bench_262144.err
timeout 900 python bench.py --envs 1048576 --steps 10 --no-cpu-baseline > $O/${TAG}_bench_1048576.json 2> $O/${TAG}_bench_1048576.err
# launch list of the bench command itself (it ran clean above): per-launch times are cold-cache
# and serialised, the kernels' SHARES are what is compared with the live numbers
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 4 --warmup 3 --mix-calls 50 --no-cpu-baseline > $O/${TAG}_ncu_launches.log 2>&1
timeout 600 python scripts/bench_replay.py > $O/${TAG}_replay_bench.json 2> $O/${TAG}_replay_bench.err
BCP_REPLAY_GRAPH=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_rp_ -c 200 --csv --log-file $O/${TAG}_replay_launches.csv python scripts/bench_replay.py --warm 190 --steps 12 > $O/${TAG}_replay_ncu.log 2>&1
BCP_LAMBDA=0 timeout 900 python bench.py --no-cpu-baseline > $O/${TAG}_bench_f32_rowspace.json 2> $O/${TAG}_bench_f32_rowspace.err
export BCP_SPLIT=1
python scripts/prof_split.py > $O/${TAG}_prof_plain.log 2>&1 && {
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve_lam -s 4 -c 1 -f -o $O/${TAG}_k_solve_lam python scripts/prof_split.py > $O/${TAG}_ncu_solve.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve_tail -s 4 -c 1 -f -o $O/${TAG}_k_solve_tail python scripts/prof_split.py > $O/${TAG}_ncu_tail.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_prep -s 7 -c 1 -f -o $O/${TAG}_k_prep python scripts/prof_split.py > $O/${TAG}_ncu_prep.log 2>&1
}
for f in $O/${TAG}_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], d.get('value'), (d.get('e2e') or {}).get('value'), d.get('kernel_ms_per_step'), (d.get('cpu_baseline') or {}).get('value'))
except Exception as e:
    print(sys.argv[1], 'unreadable', e)
PY
done
timeout 300 python scripts/bench_render.py > $O/${TAG}_render_bench.json 2> $O/${TAG}_render_bench.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render$ -s 2 -c 1 -f -o $O/${TAG}_k_render python scripts/bench_render.py --reps 3 > $O/${TAG}_ncu_render.log 2>&1
