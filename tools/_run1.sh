timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_r1_v7.json 2> gpurun_out/bench_v7.err
wc -l gpurun_out/bench_r1_v7.json
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | cut -c1-400
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 37000 -c 700 --csv --log-file gpurun_out/launches_r01e.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/ncu6.log 2>&1
