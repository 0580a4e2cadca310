export PYTHONPATH=. NSB_GEMM_DYNAMIC=0
for shape in "12288 2304 768 store 1 1" "12288 768 768 store 1 0" "12288 3072 768 store 1 1" "12288 768 3072 store 1 0" "12288 768 3072 res 1 1" "12288 768 2304 store 1 0"; do
 for cfg in "GEMM_MODE=0" "GEMM_MODE=4 GEMM_PIPE=2,0,-1" "GEMM_MODE=4 GEMM_PIPE=3,0,-1"; do
  out=$(env $cfg timeout 120 python tools/gemm_timing.py $shape 1 2>&1); echo "$cfg | $shape | $(echo "$out" | grep '50 back' | sed 's/.*back-to-back: //') | issued $(echo "$out" | grep -E '^  mma issued' | sed 's/ *mma issued *//' | cut -c1-48) | exit $(echo "$out" | grep -E '^  exit' | sed 's/ *exit *//' | cut -c1-48)"
 done
done
