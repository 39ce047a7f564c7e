set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench1.err; tail -3 gpurun_out/bench1.err
python -c "
import json; d=json.loads(open('gpurun_out/bench1.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'])"
export BCP_SPLIT=1
python scripts/prof_split.py > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_prep -s 7 -c 1 -f -o gpurun_out/prof_r1_prep4 python scripts/prof_split.py > gpurun_out/ncu3.log 2>&1
