set -x
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/t_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?" > gpurun_out/rcs.txt
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu_full.log 2>&1; echo "gpu suite rc=$?" >> gpurun_out/rcs.txt
timeout 1500 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench rc=$?" >> gpurun_out/rcs.txt
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_ref.json 2> gpurun_out/r2_bench_ref.err; echo "ref rc=$?" >> gpurun_out/rcs.txt
C4="python bench.py --config C4 --mode hutchinson --no-parity --no-cpu-baseline --steps 2 --warmup 1"
timeout 600 $C4 > gpurun_out/r2_c4_hutch.json 2> gpurun_out/r2_c4_hutch.err; rc=$?; echo "c4 hutch rc=$rc" >> gpurun_out/rcs.txt
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2_c4_fused_launches.csv $C4 > gpurun_out/ncu1.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_tc_fused -s 4 -c 2 -o gpurun_out/r2_fused $C4 > gpurun_out/ncu2.log 2>&1
fi
cat gpurun_out/rcs.txt; tail -3 gpurun_out/t_noalloc_alone.log gpurun_out/t_gpu_full.log
