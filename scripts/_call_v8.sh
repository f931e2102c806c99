set -x
mkdir -p gpurun_out
timeout 700 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_v8.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_v8.log
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_v8.json 2> gpurun_out/bench_v8.err; echo "bench rc=$?"
timeout 200 python scripts/profile_one.py > gpurun_out/profile_one_v8.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v8.csv python scripts/profile_one.py > gpurun_out/ncu_launch_v8.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 1 -c 1 -o gpurun_out/tile_v8 -f python scripts/profile_one.py > gpurun_out/ncu_tile_v8.log 2>&1
tail -3 gpurun_out/pytest_gpu_v8.log; cat gpurun_out/bench_v8.json
