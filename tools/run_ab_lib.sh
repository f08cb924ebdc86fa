#!/bin/bash
# A/B of prebuilt library variants (build/ab/libcudapt_<tag>.so) on the bench workloads (run on the GPU box):
#   gpurun -- 'TAGS="ss8 ss4" WORKLOADS="c3 c4" bash tools/run_ab_lib.sh'
cd "$(dirname "$0")/.."
cp software-path-tracer_b200/libcudapt.so /tmp/libcudapt_orig.so
for w in ${WORKLOADS:-c3}; do
  for t in $TAGS; do
    cp build/ab/libcudapt_$t.so software-path-tracer_b200/libcudapt.so
    echo "== $w $t"
    python bench.py --workload $w --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']
print(j['value'], 'Msamples/s', j['ms_per_step'], 'ms/step; e2e', j['e2e']['value'], '; trace launch_ms', r.get('launch_ms'), 'share', r.get('kernel_share_of_step'), 'packets ms', (r.get('packets') or {}).get('ms_per_step'))"
  done
done
cp /tmp/libcudapt_orig.so software-path-tracer_b200/libcudapt.so
