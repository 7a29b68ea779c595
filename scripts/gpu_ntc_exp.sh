# quick experiment loop for the narrow tensor-core kernel: its parity tests, then C2 bench lines
set -x
mkdir -p gpurun_out
rm -f gpurun_out/exp_c2*.json
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "narrow or C2 or c2 or logdensity" > gpurun_out/t_ntc.log 2>&1; echo "tests rc=$?" > gpurun_out/rcs3.txt
B="timeout 600 python bench.py --config C2 --no-cpu-baseline --steps 10 --warmup 3"
$B > gpurun_out/exp_c2_a.json 2> gpurun_out/exp_c2.err; echo "c2 rc=$?" >> gpurun_out/rcs3.txt
$B > gpurun_out/exp_c2_b.json 2>> gpurun_out/exp_c2.err
cat gpurun_out/rcs3.txt; tail -3 gpurun_out/t_ntc.log
