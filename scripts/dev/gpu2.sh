mkdir -p gpurun_out
timeout 900 python scripts/accuracy_throughput_sweep.py > gpurun_out/r1v3_accuracy_throughput.json 2> gpurun_out/r1v3_accuracy_throughput.err; tail -3 gpurun_out/r1v3_accuracy_throughput.err; cat gpurun_out/r1v3_accuracy_throughput.json
timeout 900 python scripts/accuracy_throughput_sweep.py --cart-friction 1.0 > gpurun_out/r1v3_accuracy_throughput_upstream.json 2>> gpurun_out/r1v3_accuracy_throughput.err; cat gpurun_out/r1v3_accuracy_throughput_upstream.json
