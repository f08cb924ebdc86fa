#!/bin/bash
# DRAM bytes of the per-ray traversal launches of one pass (ncu --set full), for bench.py's roofline.traffic
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for w in ${WORKLOADS:-c3 c4}; do
  spp=$([ $w = c4 ] && echo 128 || echo 64)
  n=$([ $w = c4 ] && echo 16 || echo 8)   # one pass = max_bounces closest-hit launches (the first is the packets' fall-back queue)
  ncu --set full --clock-control none -k regex:trace_kernel -c $n -o gpurun_out/traffic_$w -f \
    python tools/packet_probe.py $w $spp 400 > gpurun_out/traffic_$w.log 2>&1
  grep "render\|per packet" gpurun_out/traffic_$w.log | head -3
done
