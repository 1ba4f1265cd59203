#!/bin/bash
# kernel-row kernel variants: correctness (GPU tests) and isolated timing on the C2 / C3 shapes; tensor-pipe micro-benchmarks
out=gpurun_out/${1:-exp_kstar}; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest.log
for opt in 0 1; do
  for cfg in C2 C3; do
    SE_KSTAR_OPT=$opt timeout 300 python tools/gemm_only.py $cfg 6 16384 2>&1 | tail -1 | sed "s/^/opt=$opt /" | tee -a $out/timing.log
  done
done
for mb in 8 16 24 48 64 96; do
  SE_OZ_L2MB=$mb timeout 300 python tools/gemm_only.py C2 6 16384 2>&1 | tail -1 | tee -a $out/timing.log
done
python - <<'PY' 2>&1 | tee -a $out/timing.log
import ctypes as C, sys, os
sys.path.insert(0, os.getcwd())
from safe_exploration_b200 import _lib
lib = _lib.load()
f = C.c_double(0)
for k, what in ((3, 'N=256 1 issuer'), (8, 'N=256 2 issuers'), (9, 'N=128 1 issuer'), (10, 'N=128 2 issuers'), (5, 'N=64 1 issuer'), (7, 'N=64 2 issuers')):
    lib.se_measure_peak(k, C.byref(f)); print('burst peak kind', k, what, round(f.value / 1e12, 1), 'TOP/s')
PY
