#!/bin/bash
# ncu capture of the full-rank chain kernels (one C3 step).  Usage under gpurun: bash scripts/profile_chain.sh <tag>
TAG=${1:-r2}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python bench.py --workload c3 --profile-step --steps 1 --warmup 3 > $OUT/plain_chain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:fr_chain -s 6 -c 2 -f -o $OUT/prof_chain \
    python bench.py --workload c3 --profile-step --steps 1 --warmup 3 > $OUT/ncu_chain.log 2>&1
echo rc=$?
python scripts/ncu_summary.py $OUT/prof_chain.ncu-rep > $OUT/prof_chain.txt 2>&1
ls -la $OUT
