# stability of the default bench line: two back-to-back runs in fresh processes
set -x
mkdir -p gpurun_out
for i in 1 2; do
  timeout 1800 python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench_stab_$i.json 2> gpurun_out/r2_bench_stab_$i.err; echo "bench $i rc=$?" >> gpurun_out/rcs9.txt
done
cat gpurun_out/rcs9.txt
