#!/bin/bash
# A/B of one environment switch on the bench workloads (run on the GPU box):
#   gpurun -- 'VAR=CUDAPT_CAMERA_DIR VALUES="stored default" WORKLOADS="c3 c4" bash tools/run_ab_env.sh'
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for w in ${WORKLOADS:-c3}; do
  for v in ${VALUES:-default}; do
    if [ "$v" = default ]; then unset $VAR; else export $VAR=$v; fi
    echo "== $w $VAR=$v"
    python bench.py --workload $w --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']
print(j['value'], 'Msamples/s', j['ms_per_step'], 'ms/step; e2e', j['e2e']['value'], '; trace launch_ms', r.get('launch_ms'), 'share', r.get('kernel_share_of_step'), 'packets ms', (r.get('packets') or {}).get('ms_per_step'))"
  done
done
