set -x
mkdir -p gpurun_out
timeout 120 ./scripts/dfma_occupancy_probe > gpurun_out/dfma_probe.jsonl 2>&1
timeout 600 python scripts/variant_perf.py $VARIANTS > gpurun_out/variants.jsonl 2>&1
cat gpurun_out/dfma_probe.jsonl; cut -c1-330 gpurun_out/variants.jsonl
