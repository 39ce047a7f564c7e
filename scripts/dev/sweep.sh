for lib in libbcp.so libbcp_minb3.so; do
  for c in 2 1; do
  BCP_COHORTS=$c BCP_LIB_PATH=$PWD/cartpoleplusplus_b200/$lib python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-secondary 2>/dev/null | python -c "
import json,sys
r=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$lib cohorts $c', '%.4g'%r['value'], '%.4f'%r['ms_per_step'], r['kernel_ms_per_step'], '%.4g'%r['e2e']['value'])"
  done
done
