set -x
mkdir -p gpurun_out
CNFB200_P2P=0 timeout 150 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/t_multigpu_nccl.log 2>&1; echo "multigpu tests (NCCL exchange) rc=$?" > gpurun_out/rcs6.txt
cat gpurun_out/rcs6.txt; tail -12 gpurun_out/t_multigpu_nccl.log
