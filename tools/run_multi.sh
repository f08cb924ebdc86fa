# usage: bash tools/run_multi.sh N "workload args" [tails...]   (run under gpurun --gpus N)
N=${1:-2}
WARGS=${2:---workload c3}
shift; shift
TAILS=${@:-fused nccl}
TAG=$(echo $WARGS | tr -cd 'a-z0-9')
for tail in $TAILS; do
  echo "== N=$N $WARGS tail=$tail"
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 --tail $tail $WARGS 2>gpurun_out/multi_${N}_${TAG}_${tail}.err | tee gpurun_out/multi_${N}_${TAG}_${tail}.json | python -c "
import json,sys
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'ms/step',round(d['ms_per_step'],2),'tail',d.get('tail',{}).get('median_ms'), d['scaling'])"
  grep -i "error\|Traceback" -A5 gpurun_out/multi_${N}_${TAG}_${tail}.err | head -20
done
