set -x
nvidia-smi -L; nproc; python -c "import pybullet" 2>&1 | tail -1; ls baseline/_ref 2>&1 | head -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -8
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
