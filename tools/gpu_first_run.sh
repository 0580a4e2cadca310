#!/bin/bash
# First-contact GPU run: each test group in its own process under a timeout, logs to gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
run() { name=$1; shift; timeout 600 python -m pytest "$@" -x -q -m gpu -p no:cacheprovider > gpurun_out/$name.log 2>&1; echo "$name rc=$?" | tee -a gpurun_out/summary.txt; tail -5 gpurun_out/$name.log; }
: > gpurun_out/summary.txt
run elementwise tests/test_kernels_gpu.py -k "embedding or layernorm or cross_entropy or adamw or casts"
run gemm_simt tests/test_kernels_gpu.py -k "gemm and float32"
run attention tests/test_kernels_gpu.py -k "attention"
run gemm_tc_basic tests/test_kernels_gpu.py -k "test_gemm_kmajor_store and bfloat16"
run gemm_tc_mn tests/test_kernels_gpu.py -k "test_gemm_mn_major and bfloat16"
run gemm_tc_epi tests/test_kernels_gpu.py -k "(test_gemm_epilogues or test_gemm_wgrad_accum or tc_matches) and not float32"
cat gpurun_out/summary.txt
