#!/bin/bash
# N-GPU checks: multi-GPU parity tests, then the bench line with its strong_scaling / c4 side records.
#   gpurun --gpus N --timeout 1500 -- 'bash tools/gpu_multi.sh N [tag]'
n=${1:-2}; tag=${2:-multi$n}; out=gpurun_out/$tag; mkdir -p $out
nvidia-smi --query-gpu=index,name --format=csv > $out/gpu.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_interface.py -m gpu -q > $out/pytest_multi.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $n --steps 5 --warmup 3 > $out/bench.json 2> $out/bench.err; echo "bench rc=$?"; cut -c1-300 $out/bench.json; tail -5 $out/bench.err
python - <<PY
import json
try:
    d = json.loads(open('$out/bench.json').read().strip().splitlines()[-1])
    for k in ('value', 'ms_per_step', 'strong_scaling', 'c4', 'side_record_error'):
        print(k, json.dumps(d.get(k))[:700])
except Exception as e:
    print('no bench line', e)
PY
