#!/bin/bash
set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_ddp_gpu.py -m gpu -q > gpurun_out/pytest_ddp.log 2>&1; echo "ddp test exit $?"
tail -5 gpurun_out/pytest_ddp.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 exit $?"
tail -c 1500 gpurun_out/bench_n2.json; tail -5 gpurun_out/bench_n2.err
