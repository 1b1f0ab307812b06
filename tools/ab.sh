#!/bin/bash
# A/B timing of engine builds on one GPU box: tools/ab.sh "<label>|<lib or empty>|<env assignments>" ...
# (development probe; each case runs tools/gpu_quick.py on the bench workload)
OUT=gpurun_out/ab; mkdir -p $OUT
for spec in "$@"; do
  IFS='|' read -r label lib envs <<< "$spec"
  echo "== $label  lib=${lib:-default}  env=$envs" | tee -a $OUT/ab.log
  ( [ -n "$lib" ] && export GMR_B200_LIB=$PWD/$lib; for kv in $envs; do export $kv; done
    python tools/gpu_quick.py --reps 3 ${QUICK_ARGS} 2>&1 | grep -E "launch|rep [12]|finite|Error|error" ) | tee -a $OUT/ab.log
done
