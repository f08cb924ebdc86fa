#!/bin/bash
# configs[0] (the only workload whose CPU arm is the UNMODIFIED reference): bench lines of both arms + ncu of the megakernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
tag=${TAG:-r02}
python bench.py --workload c1 --steps 20 --warmup 3 > gpurun_out/${tag}_bench_c1.json 2> gpurun_out/${tag}_bench_c1.err || exit 1
python bench.py --impl reference --workload c1 --steps 5 --warmup 1 > gpurun_out/${tag}_ref_c1.json 2>&1
ncu --set full --clock-control none --import-source on -k regex:render_spheres -c 1 -f -o gpurun_out/${tag}_spheres_c1 \
  python bench.py --workload c1 --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_c1.log 2>&1
tail -c 400 gpurun_out/${tag}_bench_c1.json; echo; tail -c 300 gpurun_out/${tag}_ref_c1.json; ls -la gpurun_out/${tag}_spheres_c1.ncu-rep
