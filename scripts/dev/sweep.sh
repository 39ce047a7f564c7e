for cfg in "131072 BCP_LAMBDA=1 BCP_COHORTS=2" "131072 BCP_LAMBDA=1 BCP_COHORTS=4" "262144 BCP_LAMBDA=1 BCP_COHORTS=4" "262144 BCP_LAMBDA=1 BCP_COHORTS=8" "262144 BCP_LAMBDA=0 BCP_COHORTS=4" "1048576 BCP_LAMBDA=1 BCP_COHORTS=8" "1048576 BCP_LAMBDA=0 BCP_COHORTS=8" "1048576 BCP_LAMBDA=0 BCP_COHORTS=2"; do
set -- $cfg
n=$1; shift
env "$@" python bench.py --envs $n --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
r=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$cfg', '%.4g'%r['value'], '%.4f'%r['ms_per_step'], '%.4g'%r['e2e']['value'], r['config']['pipeline'])"
done
