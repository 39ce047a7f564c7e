#!/bin/bash
# One GPU-box pass that produces the artifacts summarised under profiles/ (round 2):
# tests, smoke, the bench lines (both arms, K = 20 and K = 200, fp64, large batches), the ncu launch list of the bench
# command and one --set full capture per split kernel.  Usage (repo root, on the GPU box):  bash scripts/gpu_round.sh [tag]
TAG=${1:-r2}
O=gpurun_out
mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; tail -3 $O/${TAG}_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; tail -3 $O/${TAG}_smoke.log
timeout 900 python bench.py > $O/${TAG}_bench_f32.json 2> $O/${TAG}_bench_f32.err
timeout 900 python bench.py --steps 20 --warmup 3 --no-secondary > $O/${TAG}_bench_f32_k20.json 2> $O/${TAG}_bench_f32_k20.err
timeout 900 python bench.py --impl reference --steps 20 --warmup 3 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err
timeout 900 python bench.py --dtype fp64 --steps 20 --no-cpu-baseline --no-secondary > $O/${TAG}_bench_f64.json 2> $O/${TAG}_bench_f64.err
timeout 900 python bench.py --envs 262144 --steps 40 --no-cpu-baseline --no-secondary > $O/${TAG}_bench_262144.json 2> $O/${TAG}_bench_262144.err
timeout 900 python bench.py --envs 1048576 --steps 10 --no-cpu-baseline --no-secondary > $O/${TAG}_bench_1048576.json 2> $O/${TAG}_bench_1048576.err
# launch list of the bench command itself (it ran clean above): per-launch times are cold-cache and serialised, the kernels'
# SHARES are what is compared with the live numbers.  BCP_GRAPH=0: the kernels as stream launches, not graph nodes.
export BCP_GRAPH=0
timeout 600 python bench.py --steps 4 --warmup 3 --mix-calls 50 --no-cpu-baseline --no-secondary > $O/${TAG}_plain_launches.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k_ -c 600 --csv --log-file $O/${TAG}_launches.csv \
  python bench.py --steps 4 --warmup 3 --mix-calls 50 --no-cpu-baseline --no-secondary > $O/${TAG}_ncu_launches.log 2>&1
export BCP_SPLIT=1
python scripts/prof_split.py > $O/${TAG}_prof_plain.log 2>&1 && {
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve_lam -s 8 -c 1 -f -o $O/${TAG}_k_solve_lam python scripts/prof_split.py > $O/${TAG}_ncu_solve.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve_tail -s 8 -c 1 -f -o $O/${TAG}_k_solve_tail python scripts/prof_split.py > $O/${TAG}_ncu_tail.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_prep -s 13 -c 1 -f -o $O/${TAG}_k_prep python scripts/prof_split.py > $O/${TAG}_ncu_prep.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_solve_gen_once -s 8 -c 1 -f -o $O/${TAG}_k_solve_gen_once python scripts/prof_split.py > $O/${TAG}_ncu_gen.log 2>&1
}
unset BCP_GRAPH BCP_SPLIT
for f in $O/${TAG}_bench_*.json; do python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('/')[-1], d.get('value'), (d.get('e2e') or {}).get('value'), d.get('kernel_ms_per_step'), (d.get('cpu_baseline') or {}).get('value'), (d.get('roofline') or {}).get('whole_step_frac'))
except Exception as e:
    print(sys.argv[1], 'unreadable', e)
PY
done
