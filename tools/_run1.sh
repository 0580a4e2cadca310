set -x
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
NSB_GEMM_PDL=0 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/bench_nopdl.json 2> gpurun_out/bench_nopdl.err
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/bench_pdl.json 2> gpurun_out/bench_pdl.err
NSB_GEMM_PDL=0 timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/bench_nopdl2.json 2> gpurun_out/bench_nopdl2.err
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/bench_pdl2.json 2> gpurun_out/bench_pdl2.err
grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_nopdl.json gpurun_out/bench_pdl.json gpurun_out/bench_nopdl2.json gpurun_out/bench_pdl2.json
