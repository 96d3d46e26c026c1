#!/bin/bash
# ncu evidence of round 2 (run under gpurun; every profiled command is run plain first, same arguments, `&&` before ncu).
# usage: bash scripts/ncu_evidence.sh lists|full
set -u
O=gpurun_out
NCU="ncu --clock-control none"
if [ "$1" = "lists" ]; then
  # launch lists of the bench commands: the first 2500 launches are profiled (one metric), the rest of the run is not
  for w in cfg4 cfg3 cfg2; do
    cmd="python bench.py --workload $w --steps 2 --warmup 3"
    $cmd > $O/r02_${w}_plain.log 2>&1 && \
    $NCU --metrics gpu__time_duration.sum -c 2500 --csv --log-file $O/r02_ncu_launches_${w}.csv $cmd > $O/r02_ncu_${w}.log 2>&1
    echo "$w rc=$?"
  done
else
  # one --set full capture per kernel family: launches of the SECOND pass of each target (the first pass is the warm-up)
  run() {  # name "target args" skip count regex [extra ncu flags]
    python scripts/ncu_targets.py $2 > $O/r02_$1_plain.log 2>&1 && \
    $NCU --set full ${6:-} -k regex:$5 -s $3 -c $4 -o $O/r02_ncu_full_$1 -f python scripts/ncu_targets.py $2 > $O/r02_ncu_full_$1.log 2>&1
    echo "$1 rc=$?"
    # gpurun copies back at most 64 MiB: condense the report on the box (metrics; top stall lines where the source was
    # imported) and drop it
    if [ -f $O/r02_ncu_full_$1.ncu-rep ]; then
      python scripts/ncu_summary.py $O/r02_ncu_full_$1.ncu-rep --json > $O/r02_ncu_full_$1_summary.txt 2>&1
      if [ -n "${6:-}" ]; then python scripts/ncu_top.py $O/r02_ncu_full_$1.ncu-rep 30 > $O/r02_ncu_full_$1_top.txt 2>&1; fi
      rm -f $O/r02_ncu_full_$1.ncu-rep
    fi
  }
  run ring_cfg4      "cfg4block 256"   7 4 ring_conv "--import-source on"
  run igemm_cfg2     cfg2step        114 8 igemm_conv "--import-source on"
  run halo_cfg2      cfg2step         28 4 halo_wgrad_kernel
  run ring_cfg2      cfg2step          6 3 ring_conv
  run upconv2_cfg4   "cfg4block 256"   2 2 upconv2
  run stem_cfg4      "cfg4block 256"   1 1 stem_conv
  run normapply_cfg2 cfg2step         46 2 dense_norm_apply
  run normbwd_cfg2   cfg2step         30 2 dense_norm_bwd_dx
  run reduce2_cfg2   cfg2step         61 3 dense_reduce2
  run actbwd_cfg3    cfg3step         40 2 row_act_bwd_sum
  run padfold_cfg3   cfg3step         36 2 pad_fold
  run maxpool_cfg4   "cfg4block 256"   2 1 vec_maxpool_fwd
  run pwhead_cfg4    "cfg4block 256"   1 1 pw_fprop
fi
