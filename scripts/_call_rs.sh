set -x
timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -5
for rs in "0,768" "8,768" "8,384" "16,768" "4,768" "8,1536"; do
  MPB_RS=$rs timeout 200 python scripts/phase_probe.py 8 1048576 uniform auto -1 >> gpurun_out/phase_probe_rs.jsonl 2>> gpurun_out/phase_probe_rs.err
done
MPB_RS=0,768 timeout 100 python scripts/phase_probe.py 1 1048576 uniform -1 >> gpurun_out/phase_probe_rs.jsonl 2>> gpurun_out/phase_probe_rs.err
MPB_RS=8,768 timeout 100 python scripts/phase_probe.py 1 1048576 uniform -1 >> gpurun_out/phase_probe_rs.jsonl 2>> gpurun_out/phase_probe_rs.err
tail -5 gpurun_out/phase_probe_rs.err
cut -c1-420 gpurun_out/phase_probe_rs.jsonl
