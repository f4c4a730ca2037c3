#!/bin/bash
# usage: tools/ncu_one.sh <target> <kernel-regex>   -> gpurun_out/prof_<target>.ncu-rep (second launch of the kernel)
mkdir -p gpurun_out
python tools/prof_targets.py $1 > gpurun_out/plain_$1.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$1.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$2" -s 1 -c 1 -o gpurun_out/prof_$1 -f python tools/prof_targets.py $1 > gpurun_out/ncu_$1.log 2>&1
echo "ncu $1 exit=$?"
