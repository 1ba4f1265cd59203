#!/bin/bash
# kernel-row kernel: resident threads per SM (launch bounds) 768 / 896 / 1024, isolated timing on C2 / C3 shapes
out=gpurun_out/${1:-exp_tps}; mkdir -p $out
L=safe_exploration_b200
timeout 900 python -m pytest tests -m gpu -q -x > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $out/pytest.log
cp $L/libsecem.so /tmp/libsecem_default.so
for v in default tps640 tps512; do
  if [ $v != default ]; then cp $L/libsecem_$v.so $L/libsecem.so; fi
  for cfg in C2 C3; do
    for n in 16384 8192; do
      timeout 300 python tools/gemm_only.py $cfg 6 $n 2>&1 | tail -1 | sed "s/^/$v /" | tee -a $out/timing.log
    done
  done
done
cp /tmp/libsecem_default.so $L/libsecem.so
