#!/bin/bash
# A/B of epilogue-group / staging-buffer / tile-width choices on the per-layer GEMM shapes (in-kernel traces
# through tools/gemm_timing.py).  usage (GPU box): bash tools/gemm_sweep.sh > gpurun_out/sweep.txt
export PYTHONPATH=. NSB_GEMM_DYNAMIC=0
run() { # label, env..., -- args
  label=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  out=$(env "${envs[@]}" timeout 120 python tools/gemm_timing.py "$@" 1 2>&1)
  b2b=$(echo "$out" | grep "50 back-to-back" | sed 's/.*back-to-back: //')
  ex=$(echo "$out" | grep -E "^  exit" | sed 's/ *exit *//')
  mi=$(echo "$out" | grep -E "^  mma issued" | sed 's/ *mma issued *//')
  echo "$label | $* | $b2b | mma issued $mi | exit $ex"
}
for shape in "12288 2304 768 store 1 1" "12288 768 768 store 1 0" "12288 3072 768 store 1 1" "12288 768 3072 store 1 0" "12288 768 768 res 1 1" "12288 768 3072 res 1 1" "3072 768 12288 accum 0 0" "768 768 12288 accum 0 0"; do
  run "default        " -- $shape
  run "2 groups nbuf 3" GEMM_MODE=4 GEMM_PIPE=3,0,-1 -- $shape
  run "2 groups nbuf 2" GEMM_MODE=4 GEMM_PIPE=2,0,-1 -- $shape
  run "1 group  nbuf 3" GEMM_MODE=3 GEMM_PIPE=3,0,-1 -- $shape
  run "1 group  nbuf 2" GEMM_MODE=3 GEMM_PIPE=2,0,-1 -- $shape
  run "1 grp nb4 4stage" GEMM_MODE=3 GEMM_PIPE=4,4,-1 -- $shape
  run "bn128 1cta     " GEMM_BN=128 -- $shape
done
