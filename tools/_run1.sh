for i in 1 2; do
for v in 0 1; do
NSB_OVERLAP_WGRAD=$v timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/bench_ov$v.json 2> gpurun_out/bench_ov$v.err
echo "overlap=$v $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_ov$v.json | head -1)"
done; done
