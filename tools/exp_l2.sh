#!/bin/bash
out=gpurun_out/${1:-exp_l2}; mkdir -p $out
for rep in 1 2; do
for mb in 12 16 24 32 48; do
  for cfg in C2 C3; do
    SE_OZ_L2MB=$mb timeout 300 python tools/gemm_only.py $cfg 6 16384 2>&1 | tail -1 | tee -a $out/timing.log
  done
done
done
