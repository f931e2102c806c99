#!/bin/bash
# ncu --set full capture of the second tile_kernel launch of scripts/profile_one.py; usage: r02_ncu_tile.sh <tag> [n m dtype series]
tag=$1; shift
o=gpurun_out; mkdir -p $o
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tile_kernel --launch-skip 1 -c 1 -f -o $o/${tag} \
    python scripts/profile_one.py "$@" > $o/${tag}_ncu.log 2>&1
ncu -i $o/${tag}.ncu-rep --page details --csv > $o/${tag}_ncu_details.csv 2>/dev/null
ncu -i $o/${tag}.ncu-rep --page source --csv > $o/${tag}_ncu_source.csv 2>/dev/null
tail -3 $o/${tag}_ncu.log
