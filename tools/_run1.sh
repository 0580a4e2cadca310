for i in 1 2; do
for v in 1 0; do
NSB_SPLIT_GELU=$v timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/bench_sg$v.json 2> gpurun_out/bench_sg$v.err
echo "split_gelu=$v $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_sg$v.json | head -1)"
done; done
