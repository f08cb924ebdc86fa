#!/bin/bash
# A/B of the camera-ray packets on the bench workloads (run on the GPU box: gpurun -- bash tools/run_packets.sh)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_packets.py -x -q -s 2>&1 | tail -15
for w in ${WORKLOADS:-c3}; do
  for b in ${BUDGETS:-off 160}; do
    if [ "$b" = off ]; then export CUDAPT_PRIMARY_PACKETS=0; else export CUDAPT_PRIMARY_PACKETS=1 CUDAPT_PACKET_BUDGET=$b; fi
    echo "== $w packets/budget=$b"
    python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read()); r=j['roofline']
print(j['value'], 'Msamples/s', j['ms_per_step'], 'ms/step; trace launch_ms', r.get('launch_ms'), 'share', r.get('kernel_share_of_step'))"
  done
done
