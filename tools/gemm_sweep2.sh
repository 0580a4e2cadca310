export PYTHONPATH=. NSB_GEMM_DYNAMIC=0
for cfg in "GEMM_MODE=0" "GEMM_MODE=4 GEMM_PIPE=3,0,-1" "GEMM_MODE=4 GEMM_PIPE=4,0,-1" "GEMM_MODE=3 GEMM_PIPE=6,0,-1" "GEMM_MODE=3 GEMM_PIPE=8,0,-1"; do
 for shape in "12288 768 768 res 1 1" "12288 3072 768 gelu 1 1" "12288 3072 768 dgelu 1 0"; do
  out=$(env $cfg timeout 120 python tools/gemm_timing.py $shape 1 2>&1); echo "$cfg | $shape | $(echo "$out" | grep '50 back' | sed 's/.*back-to-back: //') | exit $(echo "$out" | grep -E '^  exit' | sed 's/ *exit *//')"
 done
done
