RA=$PWD/matrix-profile-hls_b200/libmpb200_ra.so
timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
MPB200_LIB=$RA timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
python scripts/variant_perf.py default ra 2>&1 | cut -c1-270
python scripts/phase_probe.py 8 1048576 uniform auto 2>&1 | cut -c1-330
MPB200_LIB=$RA python scripts/phase_probe.py 8 1048576 uniform auto 2>&1 | cut -c1-330
MPB200_LIB=$RA python scripts/phase_probe.py 8 1048576 walk auto 2>&1 | cut -c1-330
