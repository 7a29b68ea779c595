set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu_last.log 2>&1; echo "gpu suite rc=$?" > gpurun_out/rcs19.txt
timeout 900 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_last.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rcs19.txt
timeout 600 python bench.py --config C2 --steps 10 --warmup 3 > gpurun_out/last_c2.json 2> gpurun_out/last_c2.err; echo "c2 rc=$?" >> gpurun_out/rcs19.txt
cat gpurun_out/rcs19.txt; tail -2 gpurun_out/t_gpu_last.log
