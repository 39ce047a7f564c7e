run() { timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline $2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['value'], d['kernel_ms_per_step'], d['e2e']['value'])"; }
run "base 65536" ""
BCP_PREP_BLOCK=64 run "block64 65536" ""
BCP_PREP_BLOCK=96 run "block96 65536" ""
BCP_NVCC_EXTRA="-DBCP_PREP_MINB=3" python -c "from cartpoleplusplus_b200 import build; build.build(force=True)"
grep -A2 "k_prepIfLb1" cartpoleplusplus_b200/csrc/build/bcp_f32.o.ptxas.log | tail -2 | cut -c1-200
run "minb3 65536" ""
run "minb3 262144" "--envs 262144"
