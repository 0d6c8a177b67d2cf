#!/bin/bash
# run a filtered subset of the single-GPU kernel tests with a hard timeout
set -u
OUT=${GRAFT_REPO_ROOT:-.}/gpurun_out
mkdir -p "$OUT"
timeout ${NXD_TEST_TIMEOUT:-240} python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -p no:cacheprovider -k "${NXD_TEST_FILTER:-gemm}" > "$OUT/kernel_tests.log" 2>&1; echo "kernel tests rc=$?"
tail -n 30 "$OUT/kernel_tests.log"
