run() {
  echo "== $*"
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline "$@" 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
r=d['roofline']
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'ms',round(d['ms_per_step'],2),'trace_ms/launch',round(r['launch_ms'],4),'share',round(r['kernel_share_of_step'],3))"
}
python -m pytest tests/test_gpu_mesh.py tests/test_gpu_mesh_edges.py tests/test_gpu_env.py -m gpu -x -q 2>&1 | tail -3
run --workload c3
run --workload c2
run --workload c4
