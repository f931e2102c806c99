set -x
mkdir -p gpurun_out
timeout 700 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_v7.json 2> gpurun_out/bench_v7.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_v7_ref.json 2> gpurun_out/bench_v7_ref.err; echo "ref rc=$?"
timeout 200 python scripts/profile_one.py > gpurun_out/profile_one.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v7.csv python scripts/profile_one.py > gpurun_out/ncu_launch.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 1 -c 1 -o gpurun_out/tile_v7 -f python scripts/profile_one.py > gpurun_out/ncu_tile.log 2>&1
timeout 200 python scripts/phase_probe.py 8 1048576 uniform -1 geo:6:2 > gpurun_out/phase_probe_v7.jsonl 2> gpurun_out/phase_probe_v7.err
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_v7.json; ls -la gpurun_out
