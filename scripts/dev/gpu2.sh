timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b1.json 2> gpurun_out/b1.err; tail -3 gpurun_out/b1.err
python -c "
import json; d=json.loads(open('gpurun_out/b1.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step']); print(d['roofline']); print(d['roofline_pipe'])"
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --dtype fp64 > gpurun_out/b2.json 2> gpurun_out/b2.err; tail -3 gpurun_out/b2.err
python -c "
import json; d=json.loads(open('gpurun_out/b2.json').read().strip().splitlines()[-1]); print(d['value'], d['roofline'], d['roofline_pipe'])"
