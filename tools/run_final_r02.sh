#!/bin/bash
# Final evidence of a build on ONE GPU (run on the GPU box):  gpurun --timeout 2400 -- 'TAG=r02z PART=a bash tools/run_final_r02.sh'
#   PART=a : all GPU tests, bench lines c3 (default) c1 c2 c4, reference arms c1 and c3
#   PART=b : launch lists (c3, c4), ncu --set full of the traversal launches (c3, c4), of the packet and shade kernels (c3)
#   PART=c : c5 at one GPU, the L2-band probe
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tag=${TAG:-r02z}
case ${PART:-a} in
a)
  python -m pytest tests -m gpu -q -rA --durations=10 > gpurun_out/${tag}_gputests.log 2>&1; echo "pytest exit $?" | tee -a gpurun_out/${tag}_gputests.log
  tail -4 gpurun_out/${tag}_gputests.log
  for w in c3 c1 c2 c4; do
    python bench.py --workload $w --steps ${STEPS:-10} --warmup 3 > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err
    echo "== $w rc=$?"; python -c "
import json,sys
j=json.load(open('gpurun_out/${tag}_bench_$w.json')); r=j['roofline']; c=j.get('cpu_baseline',{})
print(round(j['value'],1),'Msamples/s',round(j['ms_per_step'],3),'ms; e2e',round(j['e2e']['value'],1),'; launches',j['gpu_launches'],'; bound',r['bound'],'frac',r.get('frac'),'traffic',r.get('traffic'),'; cpu',round(c.get('value',0),2),c.get('share_of_image'))"
  done
  python bench.py --impl reference --workload c1 --steps 5 --warmup 1 > gpurun_out/${tag}_ref_c1.json 2>&1
  python bench.py --impl reference --workload c3 --steps 20 --warmup 3 > gpurun_out/${tag}_ref_c3.json 2>&1
  tail -c 300 gpurun_out/${tag}_ref_c3.json
  ;;
b)   # (gpurun copies back at most 64 MiB: the c4 capture goes in part b2)
  for w in c3 c4; do WORKLOAD=$w TAG=$tag bash tools/run_launches2.sh | tail -12; done
  WORKLOADS="c3" bash tools/run_traffic.sh
  mv gpurun_out/traffic_c3.ncu-rep gpurun_out/${tag}_traffic_c3.ncu-rep; mv gpurun_out/traffic_c3.log gpurun_out/${tag}_traffic_c3.log
  ncu --set full --clock-control none --import-source on -k regex:primary_packet -c 1 -f -o gpurun_out/${tag}_packets_c3 \
    python tools/packet_probe.py c3 64 400 > gpurun_out/ncu_pk.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:shade_kernel -c 2 -f -o gpurun_out/${tag}_shade_c3 \
    python tools/packet_probe.py c3 64 400 > gpurun_out/ncu_sh.log 2>&1
  ls -la gpurun_out/*.ncu-rep; du -sh gpurun_out
  ;;
b2)
  WORKLOADS="c4" bash tools/run_traffic.sh
  mv gpurun_out/traffic_c4.ncu-rep gpurun_out/${tag}_traffic_c4.ncu-rep; mv gpurun_out/traffic_c4.log gpurun_out/${tag}_traffic_c4.log
  ls -la gpurun_out/*.ncu-rep; du -sh gpurun_out
  ;;
c)
  python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_scale1_c5_strong_none.json 2> gpurun_out/${tag}_scale1_c5.err
  python -c "
import json; j=json.load(open('gpurun_out/${tag}_scale1_c5_strong_none.json')); print('c5 N=1', round(j['value'],1), round(j['ms_per_step'],2), round(j['e2e']['value'],1))"
  python tools/band_probe.py c3 2>&1 | tee gpurun_out/${tag}_band_probe_c3.log
  ;;
esac
