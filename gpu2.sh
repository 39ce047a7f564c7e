mkdir -p gpurun_out
run() { timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline $2 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['value'], d['kernel_ms_per_step'])"; }
for n in 131072 262144; do run "envs $n" "--envs $n"; done
for mb in 4 3 2; do
  BCP_NVCC_EXTRA="-DBCP_PREP_MINB=$mb" python -c "from cartpoleplusplus_b200 import build; build.build(force=True)"
  for pb in 128 64; do BCP_PREP_BLOCK=$pb run "minb $mb block $pb" ""; done
done
