#!/bin/bash
# BASELINE.json configs 3, 4, 5 on N GPUs (one rank per GPU, torchrun):  gpurun --gpus N -- 'bash scripts/r02_configs_multi.sh <tag> <N> [full]'
#   always: config 4 (n = 2^24, m = 1024, random walk, fp64)            -- 1 warm-up + 1 timed step per leg (a step is 34 s of one GPU)
#   full  : the multi-GPU tests, the default line (n = 2^20, m = 256, both series), config 3 (fp32, random walk), config 5 (m = 32..4096 at n = 2^22)
tag=$1; N=$2; o=gpurun_out; mkdir -p $o
run() { name=$1; shift
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) -- bench.py --gpus $N "$@" \
      > $o/${tag}_${name}_n${N}.json 2> $o/${tag}_${name}_n${N}.err
  echo "$name rc=$?"; tail -c 400 $o/${tag}_${name}_n${N}.err | tail -2; cut -c1-160 $o/${tag}_${name}_n${N}.json; }
nvidia-smi -L | wc -l
if [ "$3" = "full" ]; then
  timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $o/${tag}_pytest_multi_n${N}.log 2>&1; tail -2 $o/${tag}_pytest_multi_n${N}.log
  run default --steps 5 --warmup 3
  run config3_f32_walk --dtype f32 --data walk --steps 5 --warmup 3 --no-second-series
  run config5_msweep --n 4194304 --m-list 32,64,128,256,512,1024,2048,4096 --steps 2 --warmup 1 --no-second-series --no-cpu-baseline
fi
run config4_n2p24_m1024 --n 16777216 --m 1024 --steps 1 --warmup 1 --no-second-series --no-cpu-baseline
