set -x
python scripts/phase_probe.py 8 1048576 uniform 1,6,-1 geo:6:2 geo:6:1.5 geo:3:1.3 every:8 every:2 -1 > gpurun_out/phase_probe_uniform.jsonl 2> gpurun_out/phase_probe_uniform.err
python scripts/phase_probe.py 8 1048576 walk 1,6,-1 geo:6:2 geo:3:1.3 every:2 > gpurun_out/phase_probe_walk.jsonl 2>> gpurun_out/phase_probe_uniform.err
python scripts/phase_probe.py 1 1048576 uniform -1 > gpurun_out/phase_probe_w1.jsonl 2>> gpurun_out/phase_probe_uniform.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:tile_kernel -s 1 -c 1 -o gpurun_out/tile_v6 -f python scripts/profile_one.py > gpurun_out/ncu_tile.log 2>&1
ls -la gpurun_out
