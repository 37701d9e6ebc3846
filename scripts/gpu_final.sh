#!/bin/bash
# Round-end evidence of the current build in one GPU-box visit: parity tests, the bench line, ncu evidence
# (per-launch metrics of a step, --set full of the kernels named, bench launch list).
# usage: scripts/gpu_final.sh [kernel ...]   (default: gemm_tc_kernel)
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_T1000.json 2> gpurun_out/bench_T1000.err; echo "bench rc=$?"
cut -c1-600 gpurun_out/bench_T1000.json
timeout 900 bash scripts/ncu_evidence.sh "${@:-gemm_tc_kernel}"
