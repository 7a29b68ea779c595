set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/t_multigpu2.log 2>&1; echo "multigpu tests rc=$?" > gpurun_out/rcs6.txt
timeout 900 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rcs6.txt
timeout 900 python -m pytest tests/test_zz_no_alloc.py tests/test_golden.py -m gpu -x -q > gpurun_out/t_misc.log 2>&1; echo "misc rc=$?" >> gpurun_out/rcs6.txt
cat gpurun_out/rcs6.txt; tail -15 gpurun_out/t_multigpu2.log; cat gpurun_out/smoke.log | tail -8
