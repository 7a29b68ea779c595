# vectorised combine / error-norm kernels: the whole GPU suite (they serve the generic, per-layer and fused paths), then the C4 and default lines
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu_full5.log 2>&1; echo "gpu suite rc=$?" > gpurun_out/rcs14.txt
timeout 600 python bench.py --config C4 --mode hutchinson --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/v4_c4h.json 2> gpurun_out/v4_c4h.err; echo "c4 hutch rc=$?" >> gpurun_out/rcs14.txt
C4="python bench.py --config C4 --mode hutchinson --no-parity --no-cpu-baseline --steps 2 --warmup 1"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/v4_c4_launches.csv $C4 > gpurun_out/ncu_v4.log 2>&1
timeout 1800 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_final_n1.json 2> gpurun_out/r2_bench_final_n1.err; echo "bench rc=$?" >> gpurun_out/rcs14.txt
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_final_ref.json 2> gpurun_out/r2_bench_final_ref.err; echo "ref rc=$?" >> gpurun_out/rcs14.txt
cat gpurun_out/rcs14.txt; tail -3 gpurun_out/t_gpu_full5.log
