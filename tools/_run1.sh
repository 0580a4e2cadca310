timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
PYTHONPATH=. python tools/attn_ab.py 2>&1 | head -8
for a in "12288 768 768 store" "12288 2304 768 store"; do PYTHONPATH=. timeout 200 python tools/gemm_timing.py $a 2>&1 | sed -n 2,8p | grep -v "tile\|MMA warp"; done
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err
grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_a.json
