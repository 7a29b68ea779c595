# round-2 evidence: full GPU suite, ncu tables for the C5-shaped tensor-bound kernels, final capture of the narrow tensor kernel.
# The reports are summarised on the box (gpurun_out is limited to 64 MiB); only the C2 report travels back.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu_full4.log 2>&1; echo "gpu suite rc=$?" > gpurun_out/rcs12.txt
C5="python bench.py --config C5 --batch 131072 --no-parity --no-cpu-baseline --steps 1 --warmup 1"
timeout 900 $C5 > gpurun_out/r2_c5_b17.json 2> gpurun_out/r2_c5_b17.err; rc=$?; echo "c5 rc=$rc" >> gpurun_out/rcs12.txt
if [ $rc -eq 0 ]; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file /tmp/r2_c5_launches.csv $C5 > gpurun_out/ncu_c5a.log 2>&1
  python scripts/summarize_ncu.py launches /tmp/r2_c5_launches.csv > gpurun_out/r2_c5_launches.txt 2>&1
  timeout 1500 ncu --set full --clock-control none -k regex:"k_tc_gemm|k_tc_wgrad" -s 1500 -c 24 -o /tmp/r2_c5_kernels $C5 > gpurun_out/ncu_c5b.log 2>&1
  python scripts/summarize_ncu.py raw /tmp/r2_c5_kernels.ncu-rep > gpurun_out/r2_c5_kernels.txt 2>&1
fi
C2="python bench.py --config C2 --no-cpu-baseline --no-parity --steps 2 --warmup 1"
timeout 600 $C2 > gpurun_out/r2_c2_final.json 2> gpurun_out/r2_c2_final.err; rc=$?; echo "c2 rc=$rc" >> gpurun_out/rcs12.txt
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_c2_launches.csv $C2 > gpurun_out/ncu_c2a.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_stage_narrow_tc -s 20 -c 2 -o gpurun_out/r2_ntc_final $C2 > gpurun_out/ncu_c2b.log 2>&1
fi
du -sh gpurun_out; cat gpurun_out/rcs12.txt; tail -3 gpurun_out/t_gpu_full4.log
