timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err
grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_a.json
