#!/bin/bash
# launch list of one step of a workload (run on the GPU box, after the plain bench exited 0):
#   gpurun -- 'WORKLOAD=c3 TAG=r01d bash tools/run_launches.sh'
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
w=${WORKLOAD:-c3}; tag=${TAG:-r01d}
python bench.py --workload $w --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${w}_plain.json 2>gpurun_out/${w}_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_${w}.csv \
  python bench.py --workload $w --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
python profiles/parse_launches.py gpurun_out/${tag}_launches_${w}.csv | tee gpurun_out/${tag}_launches_${w}.summary.txt
