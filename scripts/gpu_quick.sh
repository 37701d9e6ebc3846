#!/bin/bash
# parity tests + per-kernel timing + throughput probe (no ncu)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python scripts/profile_step.py 256 bf16 > gpurun_out/profile_step_B256.log 2>&1; echo "profile rc=$?"
head -30 gpurun_out/profile_step_B256.log
timeout 300 python scripts/quick_perf.py 1 16 256 > gpurun_out/quick_perf.log 2>&1; echo "quick rc=$?"; cat gpurun_out/quick_perf.log
for extra in "$@"; do echo "== $extra"; eval "$extra"; done
