#!/bin/bash
out=gpurun_out/${1:-exp_epi}; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q -x > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $out/pytest.log
for rep in 1 2; do
  for shape in 12 0; do
    for cfg in C2 C3; do
      SE_OZ_CLUSTER=$shape timeout 300 python tools/gemm_only.py $cfg 6 16384 2>&1 | tail -1 | sed "s/^/cluster=$shape /" | tee -a $out/timing.log
    done
  done
done
SE_OZ_CLUSTER=0 SE_OZ_DBG=2 timeout 300 python tools/gemm_only.py C2 6 16384 2>&1 | tail -1 | sed "s/^/no-epilogue cluster=0 /" | tee -a $out/timing.log
bash tools/exp_env_ab.sh $1 SE_OZ_CLUSTER 12 0 2
