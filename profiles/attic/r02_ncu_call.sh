#!/bin/bash
# usage: r02_ncu_call.sh <kernel regex> <skip> <count> <out name>
mkdir -p gpurun_out
python bench.py --profile-mode --steps 2 --warmup 1 > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$1" -s $2 -c $3 -f -o gpurun_out/$4 \
    python bench.py --profile-mode --steps 2 --warmup 1 > gpurun_out/ncu_$4.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$4.log
