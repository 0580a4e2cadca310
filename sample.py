"""
Sample from a trained model: the counterpart of nanoGPT's `sample.py` on the B200-native engine
(the reference reaches it through the nanoGPT tree it clones,
/root/reference/notebooks/colab_nanoGPT_companion.ipynb:39).

    python sample.py --out_dir=out-shakespeare-char --start="ROMEO:" --num_samples=3
    python sample.py --init_from=gpt2 --start="FILE:prompt.txt"

Same keys and meaning as nanoGPT: init_from ('resume' | 'gpt2*'), out_dir, start (text, or FILE:path),
num_samples, max_new_tokens, temperature, top_k, seed, device, dtype, compile (accepted and ignored: there
is no tracing compiler on this path).  The forward runs the last-position-only lm_head (GPT.forward with
targets=None) on the tcgen05 kernels; there is no CPU path.
"""
import os
import pickle
from contextlib import nullcontext

import torch

from model import GPT, GPTConfig

# -----------------------------------------------------------------------------
init_from = 'resume'  # 'resume' (from out_dir/ckpt.pt) or a GPT-2 variant ('gpt2', 'gpt2-medium', ...)
out_dir = 'out'
start = "\n"  # prompt; "FILE:prompt.txt" reads it from a file
num_samples = 10
max_new_tokens = 500
temperature = 0.8
top_k = 200
seed = 1337
device = 'cuda'
dtype = 'bfloat16'  # 'float32' or 'bfloat16'
compile = False
exec(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'configurator.py')).read())  # --key=value overrides
# -----------------------------------------------------------------------------

if not device.startswith('cuda'):
    raise SystemExit("sample.py: device must be a CUDA device (the CPU restatement lives in oracle/ and is test-only)")
if dtype not in ('float32', 'bfloat16'):
    raise SystemExit(f"sample.py: dtype {dtype!r} is not supported (float32 or bfloat16)")
torch.manual_seed(seed)
torch.cuda.manual_seed(seed)
ctx = nullcontext() if dtype == 'float32' else torch.autocast('cuda', dtype=torch.bfloat16)

checkpoint = None
if init_from == 'resume':
    ckpt_path = os.path.join(out_dir, 'ckpt.pt')
    checkpoint = torch.load(ckpt_path, map_location='cpu')
    model = GPT(GPTConfig(**checkpoint['model_args']), verbose=False)
    state = checkpoint['model']
    for k in list(state.keys()):           # checkpoints written through torch.compile carry this prefix
        if k.startswith('_orig_mod.'):
            state[k[len('_orig_mod.'):]] = state.pop(k)
    model.load_state_dict(state)
elif init_from.startswith('gpt2'):
    model = GPT.from_pretrained(init_from, dict(dropout=0.0))
else:
    raise SystemExit(f"init_from={init_from!r}: expected 'resume' or 'gpt2*'")
model.eval()
model.to(device)

# tokenizer: the dataset's meta.pkl (character level) when the checkpoint names one, else GPT-2 BPE
encode = decode = None
if checkpoint is not None and 'config' in checkpoint and 'dataset' in checkpoint['config']:
    meta_path = os.path.join('data', checkpoint['config']['dataset'], 'meta.pkl')
    if os.path.exists(meta_path):
        with open(meta_path, 'rb') as f:
            meta = pickle.load(f)
        if 'stoi' in meta and 'itos' in meta:
            print(f"Loading meta from {meta_path}...")
            stoi, itos = meta['stoi'], meta['itos']
            encode = lambda s: [stoi[c] for c in s]
            decode = lambda l: ''.join(itos[i] for i in l)
        else:  # a token stream without a character table (the synthetic datasets): plain ids
            print(f"{meta_path} has no character table: prompt is read as space-separated token ids")
            encode = lambda s: [int(t) for t in s.split()] or [0]
            decode = lambda l: ' '.join(str(i) for i in l)
if encode is None:
    try:
        import tiktoken
        enc = tiktoken.get_encoding("gpt2")
        encode = lambda s: enc.encode(s, allowed_special={"<|endoftext|>"})
        decode = lambda l: enc.decode(l)
        print("No meta.pkl found, assuming GPT-2 encodings...")
    except Exception:  # no tokenizer available offline: fall back to raw token ids
        print("No meta.pkl and no tiktoken: prompt is read as space-separated token ids, samples are printed as ids")
        encode = lambda s: [int(t) for t in s.split()] or [0]
        decode = lambda l: ' '.join(str(i) for i in l)

if start.startswith('FILE:'):
    with open(start[5:], 'r', encoding='utf-8') as f:
        start = f.read()
start_ids = encode(start)
x = torch.tensor(start_ids, dtype=torch.long, device=device)[None, ...]

with torch.no_grad():
    with ctx:
        for k in range(num_samples):
            y = model.generate(x, max_new_tokens, temperature=temperature, top_k=top_k)
            print(decode(y[0].tolist()))
            print('---------------')
