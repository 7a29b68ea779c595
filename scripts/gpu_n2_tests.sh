set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/t_multigpu2.log 2>&1; echo "multigpu tests rc=$?" > gpurun_out/rcs6.txt
cat gpurun_out/rcs6.txt; tail -6 gpurun_out/t_multigpu2.log
