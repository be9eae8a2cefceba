#!/bin/bash
# usage: ncu_v.sh tag variant configs
tag=$1; v=$2; cfg=$3
ncu --set full --clock-control none --import-source on -k regex:'column|tiled_kernel|tc_kernel' \
    -f -o gpurun_out/${tag} python tools/quick_bench.py --configs "$cfg" --reps 1 --warmup 0 --variants $v > gpurun_out/${tag}_ncu.log 2>&1
tail -1 gpurun_out/${tag}_ncu.log
