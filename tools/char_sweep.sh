export PYTHONPATH=. NSB_GEMM_DYNAMIC=0
for shape in "16384 1152 384 store 1 1" "16384 384 384 res 1 1" "16384 1536 384 store 1 1" "16384 384 1536 res 1 1" "16384 384 384 store 1 0" "16384 1536 384 store 1 0" "16384 384 1536 store 1 0" "16384 384 1152 store 1 0"; do
 for cfg in "GEMM_MODE=0" "GEMM_MODE=4 GEMM_PIPE=3,0,-1" "GEMM_MODE=4 GEMM_PIPE=2,0,-1" "GEMM_BN=256 GEMM_MODE=0"; do
  out=$(env $cfg timeout 120 python tools/gemm_timing.py $shape 1 2>&1); echo "$cfg | $shape | $(echo "$out" | grep '50 back' | sed 's/.*back-to-back: //') | issued $(echo "$out" | grep -E '^  mma issued' | sed 's/ *mma issued *//' | cut -c1-48) | exit $(echo "$out" | grep -E '^  exit' | sed 's/ *exit *//' | cut -c1-48)"
 done
done
python - <<'PY'
import torch, time
from nanosandbox_b200.model import GPT, GPTConfig
for p in (0.0, 0.2):
    torch.manual_seed(0)
    m = GPT(GPTConfig(n_layer=6, n_head=6, n_embd=384, block_size=256, vocab_size=65, bias=False, dropout=p), verbose=False).cuda()
    opt = m.configure_optimizers(0.1, 1e-3, (0.9, 0.99), "cuda", verbose=False)
    x = torch.randint(0, 65, (64, 256), device="cuda")
    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            _, loss = m(x, x)
        loss.backward(); opt.clip_grad_norm_(1.0); opt.step(); opt.zero_grad()
    for _ in range(5): step()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(30): step()
    torch.cuda.synchronize(); print(f"char step, dropout={p}: {(time.perf_counter()-t0)/30*1e3:.2f} ms")
PY
