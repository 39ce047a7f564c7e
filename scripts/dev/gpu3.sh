set -x
mkdir -p gpurun_out
export BCP_SPLIT=1
python scripts/prof_split.py > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_prep -s 7 -c 1 -f -o gpurun_out/prof_r1_prep6 python scripts/prof_split.py > gpurun_out/ncu3.log 2>&1
