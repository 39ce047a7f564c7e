# DRAM bytes and device time of every kernel launch of two env.steps (65 536 envs, stationary mix), ncu-serialised
export BCP_SPLIT=1 BCP_GRAPH=0
python scripts/prof_split.py > gpurun_out/prof_plain.log 2>&1 && \
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:k_ -s 38 -c 36 --csv --log-file gpurun_out/r2_step_traffic.csv python scripts/prof_split.py > gpurun_out/ncu_traffic.log 2>&1
echo rc=$?
