# ncu tables for the C5-shaped (tensor-bound) forward and reverse kernels: one training step at a reduced batch
set -x
mkdir -p gpurun_out
C5="python bench.py --config C5 --batch 131072 --no-parity --no-cpu-baseline --steps 1 --warmup 1"
timeout 900 $C5 > gpurun_out/r2_c5_b17.json 2> gpurun_out/r2_c5_b17.err; rc=$?; echo "c5 rc=$rc" > gpurun_out/rcs4.txt
if [ $rc -eq 0 ]; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r2_c5_launches.csv $C5 > gpurun_out/ncu_c5a.log 2>&1
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"k_tc_gemm|k_tc_wgrad" -s 1500 -c 40 -o gpurun_out/r2_c5_kernels $C5 > gpurun_out/ncu_c5b.log 2>&1
fi
cat gpurun_out/rcs4.txt
