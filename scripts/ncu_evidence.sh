#!/bin/bash
# ncu evidence of round 2 (run under gpurun; every profiled command was run plain first, see *_plain.log).
# usage: bash scripts/ncu_evidence.sh lists|full
set -u
O=gpurun_out
NCU="ncu --clock-control none"
if [ "$1" = "lists" ]; then
  for w in cfg4 cfg3 cfg2; do
    cmd="python bench.py --workload $w --steps 2 --warmup 3"
    $cmd > $O/r02_${w}_plain.log 2>&1 && \
    $NCU --metrics gpu__time_duration.sum -c 14000 --csv --log-file $O/r02_ncu_launches_${w}.csv $cmd > $O/r02_ncu_${w}.log 2>&1
    echo "$w rc=$?"
  done
else
  # one --set full capture per kernel family, second pass of each target (first pass = warm-up)
  run() {  # name target skip count regex
    python scripts/ncu_targets.py $2 > $O/r02_$1_plain.log 2>&1 && \
    $NCU --set full --import-source on -k regex:$5 -s $3 -c $4 -o $O/r02_ncu_full_$1 -f python scripts/ncu_targets.py $2 > $O/r02_ncu_full_$1.log 2>&1
    echo "$1 rc=$?"
  }
  run ring_cfg4     "cfg4block 384"  7 7 ring_conv
  run upconv2_cfg4  "cfg4block 384"  2 2 upconv2
  run stem_cfg4     "cfg4block 384"  1 1 stem_conv
  run igemm_cfg2    cfg2step        62 6 igemm_conv
  run halo_cfg2     cfg2step        23 4 halo_wgrad_kernel
  run ring_cfg2     cfg2step         3 3 ring_conv
  run normapply_cfg2 cfg2step        5 1 dense_norm_apply
  run normbwd_cfg2   cfg2step        5 1 dense_norm_bwd_dx
  run reduce2_cfg2   cfg2step        8 2 dense_reduce2
  run actbwd_cfg3    cfg3step       12 1 row_act_bwd_sum
  run padfold_cfg3   cfg3step       12 1 pad_fold
  run loss_cfg3      cfg3step        4 2 loss_kernel
  run maxpool_cfg4   "cfg4block 384" 2 1 vec_maxpool_fwd
  run pwhead_cfg4    "cfg4block 384" 1 1 pw_fprop
fi
