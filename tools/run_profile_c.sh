#!/bin/bash
# launch list of one c3 step + full capture of the packet kernel (run on the GPU box, after the plain bench exited 0)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python bench.py --workload c3 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/c3_plain.json 2>gpurun_out/c3_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01c_launches_c3.csv \
  python bench.py --workload c3 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:primary_packet -c 1 -o gpurun_out/pkt3_c3 \
  python tools/packet_probe.py c3 64 400 > gpurun_out/ncu_pkt3.log 2>&1
tail -2 gpurun_out/ncu_pkt3.log
