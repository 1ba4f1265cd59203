#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench lines, then (only after the plain runs exited 0) the ncu launch list
# and one full capture of the two dominant kernels.  Everything lands in gpurun_out/<tag>/.
#   gpurun --timeout 1800 -- 'bash tools/gpu_validate.sh [tag] [skip-tests]'
tag=${1:-val}
out=gpurun_out/$tag
mkdir -p "$out"
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > "$out/gpu.txt" 2>&1
if [ -z "$2" ]; then
  timeout 1200 python -m pytest tests -m gpu -q --durations=10 > "$out/pytest_gpu.log" 2>&1; echo "pytest rc=$?"; tail -3 "$out/pytest_gpu.log"
  timeout 300 python __graft_entry__.py smoke > "$out/smoke.log" 2>&1; echo "smoke rc=$?"; tail -3 "$out/smoke.log"
fi
timeout 900 python bench.py --steps 10 --warmup 3 > "$out/bench_c2.json" 2> "$out/bench_c2.err"; rc=$?; echo "bench rc=$rc"; cut -c1-600 "$out/bench_c2.json"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > "$out/bench_ref.json" 2> "$out/bench_ref.err"; echo "bench ref rc=$?"; cut -c1-300 "$out/bench_ref.json"
if [ $rc -eq 0 ]; then
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file "$out/ncu_launches.csv" \
      python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-side-records > "$out/ncu_launches.log" 2>&1; echo "ncu launches rc=$?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gp_trisumsq_ozaki|gp_kstar_slices8' \
      --launch-skip 40 -c 4 -o "$out/ncu_full" -f python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-side-records \
      > "$out/ncu_full.log" 2>&1; echo "ncu full rc=$?"
  ncu -i "$out/ncu_full.ncu-rep" --page raw --csv > "$out/ncu_full_raw.csv" 2>/dev/null
fi
ls -la "$out"
