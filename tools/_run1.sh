timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_r1_v6.json 2> gpurun_out/bench_v6.err
tail -c 600 gpurun_out/bench_r1_v6.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 37000 -c 700 --csv --log-file gpurun_out/launches_r01d.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/ncu4.log 2>&1
timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --cache-control none -k regex:gemm_tc -s 17700 -c 294 --csv --log-file gpurun_out/gemm_traffic_r01d.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-torch-baseline > gpurun_out/ncu5.log 2>&1
tail -2 gpurun_out/ncu5.log
