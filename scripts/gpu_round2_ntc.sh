set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu_full2.log 2>&1; echo "gpu suite rc=$?" > gpurun_out/rcs2.txt
C2="python bench.py --config C2 --no-cpu-baseline --steps 10 --warmup 3"
timeout 600 $C2 > gpurun_out/r2_c2_ntc2.json 2> gpurun_out/r2_c2_ntc2.err; rc=$?; echo "c2 rc=$rc" >> gpurun_out/rcs2.txt
if [ $rc -eq 0 ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_stage_narrow_tc -s 20 -c 2 -o gpurun_out/r2_ntc2 python bench.py --config C2 --no-cpu-baseline --no-parity --steps 1 --warmup 1 > gpurun_out/ncu3.log 2>&1
fi
timeout 1800 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_n1_b.json 2> gpurun_out/r2_bench_n1_b.err; echo "bench rc=$?" >> gpurun_out/rcs2.txt
cat gpurun_out/rcs2.txt; tail -3 gpurun_out/t_gpu_full2.log
