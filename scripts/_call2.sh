set -x
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for sk in 0 1 2 4 8; do
  MPB_SKETCH=$sk python scripts/phase_probe.py 8 1048576 uniform -1 geo:6:2 >> gpurun_out/sketch_probe.jsonl 2>> gpurun_out/sketch_probe.err
done
for sk in 0 1 4; do
  MPB_SKETCH=$sk python scripts/phase_probe.py 1 1048576 uniform -1 >> gpurun_out/sketch_probe.jsonl 2>> gpurun_out/sketch_probe.err
  MPB_SKETCH=$sk python scripts/phase_probe.py 8 1048576 walk -1 geo:6:2 >> gpurun_out/sketch_probe.jsonl 2>> gpurun_out/sketch_probe.err
  MPB_SKETCH=$sk python scripts/phase_probe.py 1 1048576 walk -1 >> gpurun_out/sketch_probe.jsonl 2>> gpurun_out/sketch_probe.err
done
tail -3 gpurun_out/pytest_gpu.log
