#!/bin/bash
# A/B of library builds (safe_exploration_b200/libsecem_<variant>.so): isolated kernel timing on C2 / C3
#   tools/exp_lib_ab.sh TAG variant [variant ...]
out=gpurun_out/${1:-exp_lib}; shift; mkdir -p $out
L=safe_exploration_b200
cp $L/libsecem.so /tmp/libsecem_default.so
for rep in 1 2; do
for v in default "$@"; do
  if [ $v != default ]; then cp $L/libsecem_$v.so $L/libsecem.so; else cp /tmp/libsecem_default.so $L/libsecem.so; fi
  for cfg in C2 C3; do
    timeout 300 python tools/gemm_only.py $cfg 6 16384 2>&1 | tail -1 | sed "s/^/$v /" | tee -a $out/timing.log
  done
done
done
cp /tmp/libsecem_default.so $L/libsecem.so
