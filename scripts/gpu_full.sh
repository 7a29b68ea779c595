# full single-GPU regression: the whole GPU suite, smoke, the default bench and the reference arm
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu_full3.log 2>&1; echo "gpu suite rc=$?" > gpurun_out/rcs7.txt
timeout 1800 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_n1_c.json 2> gpurun_out/r2_bench_n1_c.err; echo "bench rc=$?" >> gpurun_out/rcs7.txt
timeout 900 python bench.py --config C3 --no-cpu-baseline --steps 5 --warmup 3 --tc-dbg 64 > gpurun_out/exp_c3_nopdl.json 2>> gpurun_out/r2_bench_n1_c.err; echo "c3 nopdl rc=$?" >> gpurun_out/rcs7.txt
timeout 900 python bench.py --config C3 --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/exp_c3_pdl.json 2>> gpurun_out/r2_bench_n1_c.err; echo "c3 pdl rc=$?" >> gpurun_out/rcs7.txt
cat gpurun_out/rcs7.txt; tail -3 gpurun_out/t_gpu_full3.log
