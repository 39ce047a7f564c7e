set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6
timeout 900 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench1.json 2> gpurun_out/bench1.err; tail -3 gpurun_out/bench1.err; cat gpurun_out/bench1.json
export BCP_SPLIT=1
python scripts/prof_split.py > gpurun_out/plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__inst_executed.sum,sm__inst_executed.avg.per_cycle_elapsed,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:"k_prep|k_solve" -s 5 -c 5 --csv --log-file gpurun_out/split_launches.csv python scripts/prof_split.py > gpurun_out/ncu3.log 2>&1
