#!/bin/bash
# One GPU-box visit: parity tests, per-kernel timing, throughput probe, short bench,
# ncu launch list (+ optional full capture of one kernel).  usage: scripts/gpu_check.sh [ncu_kernel_regex]
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python scripts/profile_step.py 256 bf16 > gpurun_out/profile_step_B256.log 2>&1; echo "profile rc=$?"
python scripts/quick_perf.py 1 16 256 > gpurun_out/quick_perf.log 2>&1; echo "quick rc=$?"; cat gpurun_out/quick_perf.log
python bench.py --steps 1 --warmup 3 --T 200 > gpurun_out/bench_T200.json 2> gpurun_out/bench_T200.err; echo "bench rc=$?"
NCU_CMD="python bench.py --steps 1 --warmup 3 --T 4 --no-cpu-baseline"
$NCU_CMD > gpurun_out/ncu_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 2000 -c 600 --csv --log-file gpurun_out/launches.csv $NCU_CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
if [ $# -ge 1 ]; then
  ncu --set full --clock-control none --import-source on -k regex:$1 -s 40 -c 3 -o gpurun_out/prof_$1 $NCU_CMD > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
