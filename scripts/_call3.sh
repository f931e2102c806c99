set -x
python scripts/variant_perf.py default nobatch > gpurun_out/variant_batch.jsonl 2>&1
MPB_SKETCH=2 python scripts/phase_probe.py 8 1048576 uniform -1 > gpurun_out/sketch_probe_batch.jsonl 2>&1
