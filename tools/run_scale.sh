#!/bin/bash
# Scaling lines of one build on N GPUs of one box (run on the GPU box):
#   gpurun --gpus N -- 'N=N TAG=r02 RUNS="c3:weak:samples c3:weak:tiles c3:strong:tiles c5:strong:samples" bash tools/run_scale.sh'
# Every N > 1 line carries bench.py's `parity` block (the assembled image checked against one context).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${N:-2}; tag=${TAG:-r02}; port=29600
for run in ${RUNS:-c3:weak:samples}; do
  IFS=: read w mode part <<< "$run"
  extra="${EXTRA:-}"; [ "$mode" = strong ] && [ "$w" != c5 ] && extra="$extra --strong"
  out=gpurun_out/${tag}_scale${N}_${w}_${mode}_${part}.json
  port=$((port + 1))
  if [ "$N" = 1 ]; then
    python bench.py --gpus 1 --workload $w --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline $extra > $out 2> ${out%.json}.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port bench.py \
      --gpus $N --workload $w --steps ${STEPS:-5} --warmup 3 --no-cpu-baseline --partition $part $extra > $out 2> ${out%.json}.err
  fi
  echo "== N=$N $run rc=$?"
  tail -1 $out | python -c "
import json,sys
try:
    j=json.loads(sys.stdin.read())
    p=j.get('parity') or {}
    print(round(j['value'],1), 'Msamples/s', round(j['ms_per_step'],3), 'ms/step; e2e', round(j['e2e']['value'],1), '; spp/gpu', j['config']['spp_per_gpu'], '; parity ok:', p.get('ok'), 'bit-identical-to-single:', p.get('pixels_bit_identical_to_single_context'), 'tail ms', (j.get('tail') or {}).get('median_ms'))
except Exception as e:
    print('no line', e)"
  tail -3 ${out%.json}.err | cut -c1-300
done
