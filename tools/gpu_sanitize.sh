#!/bin/bash
# compute-sanitizer (memcheck, racecheck) over __graft_entry__.smoke(): the fused-rollout solve and the tcgen05 solve.
#   gpurun --timeout 1500 -- 'bash tools/gpu_sanitize.sh [tag]'
tag=${1:-san}
out=gpurun_out/$tag
mkdir -p "$out"
timeout 300 python __graft_entry__.py smoke > "$out/smoke_plain.log" 2>&1; echo "smoke rc=$?"; tail -2 "$out/smoke_plain.log"
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --launch-timeout 0 --error-exitcode 7 --log-file "$out/sanitizer_$tool.log" \
      python __graft_entry__.py smoke > "$out/sanitizer_${tool}_stdout.log" 2>&1
  echo "$tool rc=$?"; tail -5 "$out/sanitizer_$tool.log"
done
ls -la "$out"
