set -x
mkdir -p gpurun_out
CNFB200_SLOW=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "c4_hutchinson_large or zz_tolerance" -rA > gpurun_out/t_slow.log 2>&1; echo "slow rc=$?" > gpurun_out/rcs18.txt
cat gpurun_out/rcs18.txt; tail -15 gpurun_out/t_slow.log
