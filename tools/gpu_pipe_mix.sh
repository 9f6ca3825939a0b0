#!/bin/bash
# Instruction / pipe mix of the two compositors (which pipe bounds them): candidate ncu metrics are filtered against
# what this ncu knows for the device, then collected for a few launches of render_forward / render_backward.
set -u
TAG=${1:-pipes}; O=gpurun_out/$TAG; mkdir -p $O
ncu --query-metrics --chip gb100 2>/dev/null | awk '{print $1}' > $O/all_metrics.txt || true
[ -s $O/all_metrics.txt ] || ncu --query-metrics 2>/dev/null | awk '{print $1}' > $O/all_metrics.txt
CAND="sm__inst_executed sm__inst_executed_pipe_fma sm__inst_executed_pipe_fmaheavy sm__inst_executed_pipe_fmalite sm__inst_executed_pipe_fp32 sm__inst_executed_pipe_alu sm__inst_executed_pipe_xu sm__inst_executed_pipe_lsu sm__inst_executed_pipe_fp64 sm__inst_executed_pipe_uniform sm__inst_executed_pipe_cbu sm__inst_executed_pipe_adu sm__pipe_fma_cycles_active sm__pipe_fmaheavy_cycles_active sm__pipe_fmalite_cycles_active sm__pipe_alu_cycles_active sm__pipe_xu_cycles_active sm__pipe_shared_cycles_active sm__cycles_active sm__cycles_elapsed smsp__sass_thread_inst_executed_op_ffma_pred_on smsp__sass_thread_inst_executed_op_fadd_pred_on smsp__sass_thread_inst_executed_op_fmul_pred_on smsp__sass_thread_inst_executed_op_fp32_pred_on smsp__sass_inst_executed_op_shared_ld smsp__sass_inst_executed_op_shared_st"
M=""
for c in $CAND; do
  if grep -qx "$c" $O/all_metrics.txt; then
    case $c in
      *cycles_active|*cycles_elapsed) M="$M,$c.avg,$c.sum" ;;
      *) M="$M,$c.sum" ;;
    esac
  else echo "unknown metric: $c"; fi
done
M=${M#,}
echo "metrics: $M"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$CMD > $O/plain.log 2>&1 && ncu --metrics "$M" --clock-control none -k regex:"render_" -s 6 -c 4 --csv --log-file $O/pipe_mix.csv $CMD > $O/ncu.log 2>&1
echo "ncu exit $?"
tail -3 $O/ncu.log
wc -l $O/pipe_mix.csv
