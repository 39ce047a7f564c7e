set -x
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 3 -c 1 -f -o gpurun_out/prof_r1_v2 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu3.log 2>&1
tail -3 gpurun_out/ncu3.log
