#!/bin/bash
# A/B of an environment switch on the whole C2 solve, alternating to average out box drift:
#   tools/exp_env_ab.sh TAG VAR VALUE_A VALUE_B [reps]
out=gpurun_out/${1:-exp_ab}; var=$2; a=$3; b=$4; reps=${5:-3}; mkdir -p $out
for rep in $(seq 1 $reps); do
  for v in $a $b; do
    env $var=$v timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-side-records > $out/bench_${v}_$rep.json 2> $out/bench_${v}_$rep.err
    python - <<PY | tee -a $out/summary.log
import json
d = json.loads(open('$out/bench_${v}_$rep.json').read().strip().splitlines()[-1])
r = d['roofline']
print('$var=$v rep $rep: solve %.2f ms, gemm %.4f ms, kstar %.4f ms, clock %s MHz' % (d['ms_per_step'], r['avg_launch_ms'], r['kstar_avg_launch_ms'], d['clocks']['sm_mhz']))
PY
  done
done
