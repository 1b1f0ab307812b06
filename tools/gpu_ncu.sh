#!/bin/bash
# one `ncu --set full` capture of the fused kernel on the bench workload: bash tools/gpu_ncu.sh [tag]
TAG=${1:-ncu}
OUT=gpurun_out/$TAG
mkdir -p $OUT
NCU_CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$NCU_CMD > $OUT/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/ncu_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:clip_kernel -s 1 -c 1 -o $OUT/clip_kernel $NCU_CMD > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
