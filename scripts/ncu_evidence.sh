#!/bin/bash
# ncu evidence of the current build, converted to CSV on the GPU box (the .ncu-rep files stay in /tmp there:
# gpurun returns at most 64 MiB).  Every capture runs only after the same command exited 0 without ncu.
# usage: scripts/ncu_evidence.sh [kernel ...]   (one GPU; default: every kernel of the step + the scorer)   -> gpurun_out/ev_*.csv, summarised here by scripts/ncu_evidence_summarise.sh
set -u
mkdir -p gpurun_out
export DMB200_EAGER_GRAPH=0     # the eager step as individual launches (same kernels as the graph's nodes)
STEP="python scripts/one_step.py 256 2"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__inst_issued.avg.pct_of_peak_sustained_active,l1tex__m_xbar2l1tex_read_bytes.sum
$STEP > gpurun_out/ev_plain.log 2>&1 || { echo "plain step failed"; exit 1; }
ncu --metrics $M --clock-control none --csv --log-file gpurun_out/ev_step_metrics.csv $STEP > gpurun_out/ev_step.log 2>&1; echo "step metrics rc=$?"
KERNELS="${*:-gemm_tc_kernel gn1_apply_kernel attention_mma_kernel attn_tail score_masks}"   # arguments: only these kernels
for k in $KERNELS; do
  [ "$k" = score_masks ] && continue
  ncu --set full --clock-control none --import-source on -k regex:$k -c 200 -o /tmp/ev_$k -f $STEP > gpurun_out/ev_full_$k.log 2>&1; echo "full $k rc=$?"
  ncu -i /tmp/ev_$k.ncu-rep --page raw --csv > gpurun_out/ev_full_$k.raw.csv 2>> gpurun_out/ev_full_$k.log
done
SC="python scripts/bench_scoring.py"
case " $KERNELS " in *" score_masks "*) ;; *) SC=false ;; esac
$SC > gpurun_out/ev_scoring_plain.log 2>&1 && {
  ncu --set full --clock-control none --import-source on -k regex:score_masks -c 2 -o /tmp/ev_score -f $SC > gpurun_out/ev_full_score.log 2>&1; echo "full score rc=$?"
  ncu -i /tmp/ev_score.ncu-rep --page raw --csv > gpurun_out/ev_full_score_masks.raw.csv 2>> gpurun_out/ev_full_score.log
}
# launch list of the bench command itself (gpu__time_duration only), after a plain run of the same command
NCU_CMD="python bench.py --steps 1 --warmup 3 --T 4 --no-cpu-baseline --no-extras"
$NCU_CMD > gpurun_out/ev_bench_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/ev_bench_launches.csv $NCU_CMD > gpurun_out/ev_bench_ncu.log 2>&1
echo "bench launches rc=$?"
ls -la gpurun_out/ev_* | awk '{print $5, $9}'
