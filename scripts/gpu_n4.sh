# 4-GPU sanity of the default bench exactly as the driver launches it (plus the multi-GPU tests)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/t_multigpu4.log 2>&1; echo "multigpu tests rc=$?" > gpurun_out/rcs13.txt
timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29527 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/r2_bench_n4.json 2> gpurun_out/r2_bench_n4.err; echo "bench n4 rc=$?" >> gpurun_out/rcs13.txt
cat gpurun_out/rcs13.txt; tail -3 gpurun_out/t_multigpu4.log; tail -c 600 gpurun_out/r2_bench_n4.err
