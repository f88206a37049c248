#!/bin/bash
# the -DQMMX_DEBUG_BOUNDS build (every shared-memory access of the lean / tile kernels checked, a violation traps) under the
# GPU parity tests that reach every kernel family at sizes the checks can afford
mkdir -p gpurun_out
QMMX_LIB=$PWD/qmmx_monolithic_monte_carlo_b200/libqmmx_mc_dbg.so timeout 1500 python -m pytest tests/test_gpu_parity.py -m gpu -q \
  -k "not full and not lean_kernels_equal and not wide_level_sets_first" 2>&1 | tail -6 | tee gpurun_out/bounds_check_build.log
