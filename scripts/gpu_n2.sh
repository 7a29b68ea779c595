# 2-GPU regression: the multi-GPU tests, then the default bench under torchrun exactly as the driver launches it
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q > gpurun_out/t_multigpu.log 2>&1; echo "multigpu tests rc=$?" > gpurun_out/rcs5.txt
timeout 2400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench n2 rc=$?" >> gpurun_out/rcs5.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29518 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r2_bench_ref_n2.json 2> gpurun_out/r2_bench_ref_n2.err; echo "ref n2 rc=$?" >> gpurun_out/rcs5.txt
cat gpurun_out/rcs5.txt; tail -3 gpurun_out/t_multigpu.log; tail -c 600 gpurun_out/r2_bench_n2.err
