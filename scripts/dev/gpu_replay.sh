set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_replay.py -m gpu -x -q 2>&1 | tail -25
timeout 600 python scripts/bench_replay.py > gpurun_out/replay_bench.json 2> gpurun_out/replay_bench.err; tail -3 gpurun_out/replay_bench.err; cat gpurun_out/replay_bench.json
timeout 200 python scripts/dev/replay_micro.py
BCP_REPLAY_GRAPH=0 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:k_rp_ -c 200 --csv --log-file gpurun_out/replay_launches.csv python scripts/bench_replay.py --warm 190 --steps 12 > gpurun_out/replay_ncu.log 2>&1
tail -2 gpurun_out/replay_ncu.log
