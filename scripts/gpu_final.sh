# final state check: what the driver runs at round end (GPU suite, smoke, default bench, reference arm)
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu_final.log 2>&1; echo "gpu suite rc=$?" > gpurun_out/rcs16.txt
timeout 900 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?" >> gpurun_out/rcs16.txt
timeout 1800 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_final2_n1.json 2> gpurun_out/r2_bench_final2_n1.err; echo "bench rc=$?" >> gpurun_out/rcs16.txt
timeout 600 python bench.py > gpurun_out/r2_bench_noflags.json 2> gpurun_out/r2_bench_noflags.err; echo "bench noflags rc=$?" >> gpurun_out/rcs16.txt
cat gpurun_out/rcs16.txt; tail -3 gpurun_out/t_gpu_final.log; tail -4 gpurun_out/smoke_final.log
