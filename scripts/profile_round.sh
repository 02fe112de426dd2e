#!/bin/bash
# Run on the GPU box (under gpurun):  bash scripts/profile_round.sh <tag> [workloads...]
# For every workload: a plain run, then the ncu launch list of the device-resident step, then one
# `ncu --set full` capture of the dominant kernel(s).  Everything lands in gpurun_out/<tag>/.
# Numbers printed under ncu are never bench values (B200_PROFILING.md).
set -u
TAG=${1:-r1}
shift || true
WLS=${@:-"c4 c5 c3 c2"}
OUT=gpurun_out/$TAG
mkdir -p $OUT
declare -A KREGEX=( [c4]="mf_step_kernel" [c1]="mf_step_kernel" [c5]="filter_kernel|normalise_kernel|divergence_blocks"
                    [c3]="gemm_tf32x3_kernel" [c3chol]="gemm_tf32x3_kernel" [c2]="reg_fused_kernel" [n3]="psis_columns_kernel" [n4r]="rhat_halfchain_kernel" )
for w in $WLS; do
  echo "== $w: launch list"
  python bench.py --workload $w --profile-step --steps 2 --warmup 3 > $OUT/plain_step_$w.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/launches_$w.csv \
      python bench.py --workload $w --profile-step --steps 2 --warmup 3 > $OUT/ncu_step_$w.log 2>&1
  echo "   rc=$?"
  python scripts/launch_summary.py $OUT/launches_$w.csv > $OUT/launches_$w.txt 2>&1
  [ "${FULL:-1}" = 0 ] && continue
  echo "== $w: full capture of ${KREGEX[$w]}"
  SKIP=2; [ "$w" = c5 ] && SKIP=3
  python bench.py --workload $w --profile-kernel-only --steps 2 --warmup 3 > $OUT/plain_kernel_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k "regex:${KREGEX[$w]}" -s $SKIP -c 3 -f -o $OUT/prof_$w \
      python bench.py --workload $w --profile-kernel-only --steps 2 --warmup 3 > $OUT/ncu_kernel_$w.log 2>&1
  echo "   rc=$?"
  # the text summary is what gets committed; KEEP_REP=1 keeps the report for source-level reading (gpurun_out/ is
  # capped at 64 MiB per call)
  python scripts/ncu_summary.py $OUT/prof_$w.ncu-rep > $OUT/prof_$w.txt 2>&1
  [ "${KEEP_REP:-0}" = 1 ] || rm -f $OUT/prof_$w.ncu-rep
done
rm -f $OUT/*.csv
ls -la $OUT
