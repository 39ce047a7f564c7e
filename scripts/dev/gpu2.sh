run() { timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline $2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['value'], d['kernel_ms_per_step'])"; }
for n in 65536 98304 131072 196608 524288; do for w in 0 1; do BCP_PREP_WIDE=$w run "envs $n wide $w" "--envs $n"; done; done
