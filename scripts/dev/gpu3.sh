mkdir -p gpurun_out
export BCP_SPLIT=1
python scripts/prof_split.py > gpurun_out/plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,sm__inst_executed.avg.per_cycle_elapsed,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__warps_launched.sum --clock-control none -k regex:"k_prep|k_solve" -s 7 -c 7 --csv --log-file gpurun_out/split_launches.csv python scripts/prof_split.py > gpurun_out/ncu3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_solve -s 5 -c 1 -f -o gpurun_out/prof_r1_solve_p1 python scripts/prof_split.py > gpurun_out/ncu4.log 2>&1
