#!/bin/bash
# launch list of one step of a workload with lane / issue counters (run on the GPU box, after the plain bench exited 0):
#   gpurun -- 'WORKLOAD=c3 TAG=r02 bash tools/run_launches2.sh'
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
w=${WORKLOAD:-c3}; tag=${TAG:-r02}
python bench.py --workload $w --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_${w}_plain.json 2>gpurun_out/${tag}_${w}_plain.err || exit 1
ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,sm__inst_issued.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed \
  --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches_${w}.csv \
  python bench.py --workload $w --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
python profiles/parse_launches.py gpurun_out/${tag}_launches_${w}.csv | tee gpurun_out/${tag}_launches_${w}.summary.txt
