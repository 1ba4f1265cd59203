#!/bin/bash
# kernel-row tail balancing (training set cut over blockIdx.y for large batches): GPU tests, isolated timing per cut, bench
out=gpurun_out/${1:-exp_tail}; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q -x > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest.log
for cfg in C2 C3; do
  for tail in 0 1 2 4 8 16; do
    SE_KSTAR_TAIL=$tail timeout 300 python tools/gemm_only.py $cfg 6 16384 2>&1 | tail -1 | sed "s/^/tail=$tail /" | tee -a $out/timing.log
  done
done
for n in 8192 12288 20480 32768; do
  for tail in 0 1; do
    SE_KSTAR_TAIL=$tail timeout 300 python tools/gemm_only.py C2 6 $n 2>&1 | tail -1 | sed "s/^/tail=$tail /" | tee -a $out/timing.log
  done
done
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-side-records > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; cut -c1-400 $out/bench.json
