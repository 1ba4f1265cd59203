#!/bin/bash
out=gpurun_out/${1:-exp_factor}; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_train.py -m gpu -q -x > $out/pytest_train.log 2>&1; echo "pytest train rc=$?"; tail -15 $out/pytest_train.log
timeout 900 python -m pytest tests -m gpu -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -8 $out/pytest.log
timeout 600 python tools/update_probe.py > $out/update_probe.log 2>&1; tail -12 $out/update_probe.log
