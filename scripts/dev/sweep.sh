for v in "BCP_PREP2=1" "BCP_PREP2=0" "BCP_PREP2=1 BCP_COHORTS=1" "BCP_PREP2=0 BCP_COHORTS=1" "BCP_PREP2=1 BCP_COHORTS=4"; do
  env $v python bench.py --steps 40 --warmup 5 --no-cpu-baseline --no-secondary 2>/dev/null | python -c "
import json,sys
r=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$v', '%.4g'%r['value'], '%.4f'%r['ms_per_step'], r['kernel_ms_per_step'], 'e2e %.4g'%r['e2e']['value'])"
done
