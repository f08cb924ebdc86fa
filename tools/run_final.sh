# final single-GPU bench lines of the round (JSON kept under gpurun_out/, copied to profiles/ afterwards)
for w in c1 c2 c3 c4; do
  python bench.py --workload $w --steps 5 --warmup 3 2>gpurun_out/final_$w.err > gpurun_out/final_$w.json; tail -c 600 gpurun_out/final_$w.json; echo
done
python bench.py --impl reference --workload c1 --steps 3 --warmup 1 > gpurun_out/final_ref_c1.json 2>/dev/null
python bench.py --impl reference --workload c3 --steps 2 --warmup 1 > gpurun_out/final_ref_c3.json 2>/dev/null
cat gpurun_out/final_ref_c1.json gpurun_out/final_ref_c3.json | cut -c1-300
