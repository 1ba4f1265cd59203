#!/bin/bash
# where does the variance GEMM's time go: operand-traffic knobs of the kernel (SE_OZ_DBG bits: 1 no loads, 2 no epilogue,
# 4 no candidate digits, 8 no L^-1 digits) on the C2 shape, isolated launches
out=gpurun_out/${1:-exp_gemm}; mkdir -p $out
for dbg in 0 1 2 3 4 8 6 10; do
  SE_OZ_DBG=$dbg timeout 300 python tools/gemm_only.py C2 6 16384 2>&1 | tail -1 | tee -a $out/timing.log
done
