#!/bin/bash
# usage: prof1.sh <tag> ; env passes through
OUT=gpurun_out/$1; mkdir -p $OUT
python bench.py --workload c2 --profile-kernel-only --steps 2 --warmup 3 > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:gemm_tf32x3_kernel" -s 2 -c 2 -f -o $OUT/prof python bench.py --workload c2 --profile-kernel-only --steps 2 --warmup 3 > $OUT/ncu.log 2>&1
echo rc=$?
