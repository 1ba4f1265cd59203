#!/bin/bash
out=gpurun_out/${1:-exp_gemm2}; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_tc.py -m gpu -q -x > $out/pytest_tc.log 2>&1; echo "pytest rc=$?"; tail -2 $out/pytest_tc.log
for rep in 1 2; do
for mb in 32 24; do
  SE_OZ_L2MB=$mb timeout 300 python tools/gemm_only.py C2 6 16384 2>&1 | tail -1 | tee -a $out/timing.log
done
done
for d in 5 7 8; do timeout 300 python tools/gemm_only.py C2 $d 16384 2>&1 | tail -1 | tee -a $out/timing.log; done
timeout 300 python tools/gemm_only.py C3 6 16384 2>&1 | tail -1 | tee -a $out/timing.log
