timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_r1_v8.json 2> gpurun_out/bench_v8.err
wc -l gpurun_out/bench_r1_v8.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 37000 -c 700 --csv --log-file gpurun_out/launches_r01f.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/ncu8.log 2>&1
