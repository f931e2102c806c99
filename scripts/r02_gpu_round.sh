#!/bin/bash
# One gpurun call of round 2: GPU tests, the default bench line + the reference arm, the ncu launch list and one
# `ncu --set full` capture of the tile kernel on the north_star workload (random walk, n = 2^20, m = 256, fp64 default mode).
# Usage: gpurun --timeout 1800 -- 'bash scripts/r02_gpu_round.sh [tag]'.  Everything lands in gpurun_out/<tag>_*.
tag=${1:-r02}
o=gpurun_out
mkdir -p $o
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $o/${tag}_gpu.txt
( time timeout 1500 python -m pytest tests -m gpu -x -q ) > $o/${tag}_pytest.log 2>&1
echo "pytest rc=$?" >> $o/${tag}_pytest.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $o/${tag}_pytest_noalloc_alone.log 2>&1
timeout 300 python __graft_entry__.py smoke > $o/${tag}_smoke.log 2>&1
timeout 600 python bench.py --steps 5 --warmup 3 > $o/${tag}_bench.json 2> $o/${tag}_bench.err
timeout 300 python bench.py --impl reference --steps 1 --warmup 1 > $o/${tag}_bench_reference_arm.json 2>> $o/${tag}_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_launches_walk.csv \
    python scripts/profile_one.py 1048576 256 f64 walk > $o/${tag}_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tile_kernel --launch-skip 1 -c 1 -f -o $o/${tag}_tile_walk \
    python scripts/profile_one.py 1048576 256 f64 walk > $o/${tag}_ncu_full.log 2>&1
ncu -i $o/${tag}_tile_walk.ncu-rep --page details --csv > $o/${tag}_tile_kernel_walk_ncu_details.csv 2>/dev/null
tail -3 $o/${tag}_pytest.log; tail -2 $o/${tag}_pytest_noalloc_alone.log; cut -c1-600 $o/${tag}_bench.json
