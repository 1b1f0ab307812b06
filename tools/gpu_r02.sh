#!/bin/bash
# One GPU-box visit of round 2.  usage (repo root, under gpurun): bash tools/gpu_r02.sh TAG [steps...]
#   steps: test smoke bench ref ops full launches   (default: all)
TAG=${1:-r02}; shift
STEPS=${@:-test smoke bench ops full launches}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/gpu.txt 2>&1
nproc >> $OUT/gpu.txt
NCU_CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras"
OPS=smsp__sass_thread_inst_executed_op_dfma_pred_on.sum,smsp__sass_thread_inst_executed_op_dmul_pred_on.sum,smsp__sass_thread_inst_executed_op_dadd_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,gpu__time_duration.sum,sm__warps_active.avg.pct_of_peak_sustained_active
for s in $STEPS; do
  case $s in
    test) python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu.log; tail -3 $OUT/pytest_gpu.log ;;
    smoke) python __graft_entry__.py --smoke > $OUT/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/smoke.log ;;
    bench) python bench.py > $OUT/bench.json 2> $OUT/bench.err; echo "bench rc=$?"; tail -c 1500 $OUT/bench.json ;;
    ref) python bench.py --impl reference > $OUT/bench_reference.json 2> $OUT/bench_reference.err; echo "ref rc=$?" ;;
    ops) $NCU_CMD > $OUT/ncu_plain.log 2>&1 &&
         ncu --metrics $OPS --clock-control none -k regex:clip_kernel -s 1 -c 1 --csv --log-file $OUT/clip_kernel_ops.csv $NCU_CMD > $OUT/ncu_ops.log 2>&1
         echo "ncu ops rc=$?" ;;
    full) $NCU_CMD > $OUT/ncu_plain.log 2>&1 &&
          ncu --set full --clock-control none --import-source on -k regex:clip_kernel -s 1 -c 1 -f -o $OUT/clip_kernel $NCU_CMD > $OUT/ncu_full.log 2>&1
          echo "ncu full rc=$?" ;;
    launches) $NCU_CMD > $OUT/ncu_plain.log 2>&1 &&
          ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $OUT/launches.csv $NCU_CMD > $OUT/ncu_launches.log 2>&1
          echo "ncu launches rc=$?" ;;
  esac
done
ls -la $OUT
