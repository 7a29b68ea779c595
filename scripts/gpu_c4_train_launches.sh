set -x
mkdir -p gpurun_out
C4="python bench.py --config C4 --strong --no-parity --no-cpu-baseline --steps 1 --warmup 1"
timeout 600 $C4 > gpurun_out/c4_train.json 2> gpurun_out/c4_train.err; echo "c4 train rc=$?" > gpurun_out/rcs17.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file /tmp/c4_train_launches.csv $C4 > gpurun_out/ncu_c4t.log 2>&1
python scripts/summarize_ncu.py launches /tmp/c4_train_launches.csv > gpurun_out/r2_c4_train_launches.txt 2>&1
cat gpurun_out/rcs17.txt; head -30 gpurun_out/r2_c4_train_launches.txt
