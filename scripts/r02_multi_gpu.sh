#!/bin/bash
# Multi-GPU gpurun call: usage  gpurun --gpus N -- 'bash scripts/r02_multi_gpu.sh <tag> <N> [tests]'
# runs the multi-GPU tests (optional) and the default bench line on N GPUs (torchrun, one rank per GPU)
tag=$1; N=$2; o=gpurun_out; mkdir -p $o
nvidia-smi -L > $o/${tag}_gpus.txt
if [ "$3" = "tests" ]; then timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $o/${tag}_pytest_multi.log 2>&1; tail -3 $o/${tag}_pytest_multi.log; fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3 \
    > $o/${tag}_bench_n${N}.json 2> $o/${tag}_bench_n${N}.err
tail -2 $o/${tag}_bench_n${N}.err; cut -c1-300 $o/${tag}_bench_n${N}.json
