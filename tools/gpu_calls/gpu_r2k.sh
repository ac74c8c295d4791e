#!/bin/bash
# 2 GPUs: multi-rank sharded-search tests, then the default bench under torchrun (c3 headline + other_configs.c5)
mkdir -p gpurun_out/r2k
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/r2k/smi.txt
timeout 600 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_parity.py -x -q -m gpu -k "shard or search or scatter or classifier" > gpurun_out/r2k/pytest_sharded.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/r2k/pytest_sharded.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/r2k/bench_2gpu.json 2> gpurun_out/r2k/bench_2gpu.err; echo "bench2 exit $?"
tail -c 2500 gpurun_out/r2k/bench_2gpu.json; tail -3 gpurun_out/r2k/bench_2gpu.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --workload c5 --exchange nccl --steps 5 --warmup 3 > gpurun_out/r2k/bench_c5_2gpu_nccl.json 2> gpurun_out/r2k/bench_c5_2gpu_nccl.err; echo "c5 nccl exit $?"
tail -c 1200 gpurun_out/r2k/bench_c5_2gpu_nccl.json
