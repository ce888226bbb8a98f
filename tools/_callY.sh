timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
bash tools/profile_round.sh r02b 2>&1 | tail -5
ls gpurun_out | tail -12
