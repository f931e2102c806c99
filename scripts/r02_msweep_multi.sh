#!/bin/bash
# BASELINE.json config 5 (m = 32..4096 at n = 2^22, random walk, fp64) on N GPUs: gpurun --gpus N -- 'bash scripts/r02_msweep_multi.sh <tag> <N>'
tag=$1; N=$2; o=gpurun_out; mkdir -p $o
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) -- bench.py --gpus $N \
   --n 4194304 --m-list 32,64,128,256,512,1024,2048,4096 --steps 1 --warmup 1 --no-second-series --no-cpu-baseline > $o/${tag}_config5_msweep_n${N}.json 2> $o/${tag}_config5_msweep_n${N}.err
echo rc=$?; tail -c 300 $o/${tag}_config5_msweep_n${N}.err; cut -c1-150 $o/${tag}_config5_msweep_n${N}.json
