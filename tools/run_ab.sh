run() {
  echo "== $*"
  env "$@" python bench.py --workload c3 --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('value',round(d['value'],1),'ms',round(d['ms_per_step'],2),'trace_ms/launch',round(r['launch_ms'],4),'share',round(r['kernel_share_of_step'],3),'Nn',round(r['Nn'],3),'Nt',round(r['Nt'],3))"
}
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run A=1
run CUDAPT_TRACE_BLOCKS=9
run CUDAPT_TRACE_BLOCKS=6
