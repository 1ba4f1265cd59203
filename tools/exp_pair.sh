#!/bin/bash
# cluster variance GEMM (digit planes multicast inside a CA x CB thread-block cluster): parity tests, isolated timing
out=gpurun_out/${1:-exp_pair}; mkdir -p $out
for shape in 22 41 14; do
  SE_OZ_CLUSTER=$shape timeout 600 python -m pytest tests/test_gpu_tc.py tests/test_gpu_configs.py -m gpu -q -x > $out/pytest_$shape.log 2>&1; echo "pytest(cluster $shape) rc=$?"; tail -1 $out/pytest_$shape.log
done
for rep in 1 2; do
  for shape in 0 12 21 22 14 41; do
    for cfg in C2 C3; do
      SE_OZ_CLUSTER=$shape timeout 300 python tools/gemm_only.py $cfg 6 16384 2>&1 | tail -1 | sed "s/^/cluster=$shape /" | tee -a $out/timing.log
    done
  done
done
