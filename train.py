"""Training driver with nanoGPT's ``train.py`` contract, on the B200-native step.

Single GPU:
    python train.py config/train_shakespeare_char.py --dropout=0.0 --compile=False
DDP on one node / one Pod (/root/reference/README.md:7, 55-60):
    torchrun --standalone --nproc_per_node=8 train.py config/train_gpt2.py
DDP across Pods (StatefulSet, /root/reference/README.md:8, 62-72; deploy/entrypoint.sh sets NODE_RANK):
    torchrun --nnodes=8 --node_rank=$NODE_RANK --nproc_per_node=1 \
             --master_addr=$MASTER_ADDR --master_port=$MASTER_PORT train.py config/train_gpt2_medium.py

Same globals, same ``config.py --key=value`` overrides, same RANK / LOCAL_RANK / WORLD_SIZE
handling, same log line as nanoGPT.  Underneath: nanosandbox_b200.GPT (one
fused forward/backward on sm_100a kernels), FusedAdamW (clip folded into the update) and our
bucketed NCCL all-reduce instead of torch DDP.  CUDA only: ``--device=cpu`` is refused.
"""
import math
import os
import pickle
import time
from contextlib import nullcontext

import numpy as np
import torch
from torch.distributed import destroy_process_group, init_process_group

from nanosandbox_b200.ddp import DistributedDataParallel as DDP
from nanosandbox_b200.model import GPT, GPTConfig

# -----------------------------------------------------------------------------
# default config values: GPT-2 (124M) on OpenWebText-shaped data
# I/O
out_dir = 'out'
eval_interval = 2000
log_interval = 1
eval_iters = 200
eval_only = False
always_save_checkpoint = True
init_from = 'scratch'  # 'scratch', 'resume' or 'gpt2*'
# wandb logging (accepted for CLI compatibility; not available offline)
wandb_log = False
wandb_project = 'owt'
wandb_run_name = 'gpt2'
# data
dataset = 'openwebtext'
gradient_accumulation_steps = 5 * 8
batch_size = 12
block_size = 1024
# model
n_layer = 12
n_head = 12
n_embd = 768
dropout = 0.0
bias = False
# adamw optimizer
learning_rate = 6e-4
max_iters = 600000
weight_decay = 1e-1
beta1 = 0.9
beta2 = 0.95
grad_clip = 1.0
# learning rate decay settings
decay_lr = True
warmup_iters = 2000
lr_decay_iters = 600000
min_lr = 6e-5
# DDP settings
backend = 'nccl'
# system
device = 'cuda'
dtype = 'bfloat16'
compile = True  # accepted and ignored: there is no tracing compiler on this path
# ---- additions of this implementation (all optional) ----
synthetic = True         # generate a seeded synthetic token stream when data/<dataset>/train.bin is absent
synthetic_kind = 'uniform'  # 'uniform' (throughput) or 'markov' (learnable, loss-curve checks)
synthetic_vocab = 0      # 0: 65 for *_char datasets, else 50304
bucket_dtype = 'bf16'    # gradient all-reduce payload: 'bf16' (cast+scale fused in the pack kernel) or 'fp32'
peak_flops = 2.25e15     # B200 dense bf16 peak used for the second MFU figure in the log line
tensorboard_log = False  # write loss / lr / tokens-per-second scalars (the reference mounts /data/runs for TensorBoard)
tensorboard_dir = 'runs' # e.g. /data/runs/<job> in the k8s Jobs (/root/reference/README.md:83-87)
# -----------------------------------------------------------------------------
config_keys = [k for k, v in globals().items() if not k.startswith('_') and isinstance(v, (int, float, bool, str))]
exec(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), 'configurator.py')).read())
config = {k: globals()[k] for k in config_keys}
# -----------------------------------------------------------------------------

ddp = int(os.environ.get('RANK', -1)) != -1
if ddp:
    ddp_rank = int(os.environ['RANK'])
    ddp_local_rank = int(os.environ['LOCAL_RANK'])
    ddp_world_size = int(os.environ['WORLD_SIZE'])
    device = f'cuda:{ddp_local_rank}'
    torch.cuda.set_device(device)
    init_process_group(backend=backend, device_id=torch.device(device) if backend == 'nccl' else None)
    master_process = ddp_rank == 0
    seed_offset = ddp_rank
    assert gradient_accumulation_steps % ddp_world_size == 0, \
        "gradient_accumulation_steps must be a multiple of WORLD_SIZE"
    gradient_accumulation_steps //= ddp_world_size
else:
    master_process = True
    seed_offset = 0
    ddp_world_size = 1
tokens_per_iter = gradient_accumulation_steps * ddp_world_size * batch_size * block_size
if master_process:
    print(f"tokens per iteration will be: {tokens_per_iter:,}")
    os.makedirs(out_dir, exist_ok=True)

device_type = 'cuda' if 'cuda' in device else 'cpu'
if device_type != 'cuda' or not torch.cuda.is_available():
    raise SystemExit("train.py: this implementation runs on CUDA (sm_100a) only; there is no CPU "
                     "fallback. The CPU restatement used for parity lives in oracle/.")
if dtype not in ('bfloat16', 'float32'):
    raise SystemExit(f"train.py: dtype={dtype!r} is not supported (bfloat16 or float32)")
if compile and master_process:
    print("note: --compile is accepted for compatibility and ignored (hand-written kernels, no torch.compile)")
torch.manual_seed(1337 + seed_offset)
ctx = nullcontext() if dtype == 'float32' else torch.amp.autocast(device_type='cuda', dtype=torch.bfloat16)

# -----------------------------------------------------------------------------
# data: train.bin / val.bin of uint16 tokens (+ meta.pkl), as nanoGPT's prepare.py leaves them
data_dir = os.path.join('data', dataset)
# Whether the data has to be synthesised is decided collectively (any rank without train.bin -> yes), so all
# ranks take the same collectives in the same order.  One writer per node (LOCAL_RANK 0: in the 8-Pod
# StatefulSet every Pod has its own filesystem) writes the seeded stream (identical everywhere) to temporary
# names and renames them into place, meta.pkl last, so no reader ever sees a partly written file.
need_synth = not os.path.exists(os.path.join(data_dir, 'train.bin'))
if ddp:
    flag = torch.tensor([int(need_synth)], device=device)
    torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MAX)
    need_synth = bool(flag.item())
if need_synth:
    if not synthetic:
        raise SystemExit(f"{data_dir}/train.bin not found and --synthetic=False")
    from nanosandbox_b200.data import write_dataset
    vocab = synthetic_vocab or (65 if dataset.endswith('_char') else 50304)
    if (not ddp) or ddp_local_rank == 0:
        if master_process:
            print(f"{data_dir}/train.bin not found: writing a seeded synthetic {synthetic_kind} stream, vocab {vocab}")
        if not os.path.exists(os.path.join(data_dir, 'train.bin')):
            write_dataset(data_dir, vocab, kind=synthetic_kind)
if ddp:
    torch.distributed.barrier()


def get_batch(split):
    # a fresh memmap every batch, as nanoGPT does (avoids a numpy memmap leak)
    data = np.memmap(os.path.join(data_dir, 'train.bin' if split == 'train' else 'val.bin'), dtype=np.uint16, mode='r')
    ix = torch.randint(len(data) - block_size, (batch_size,))
    x = torch.stack([torch.from_numpy((data[i:i + block_size]).astype(np.int64)) for i in ix])
    y = torch.stack([torch.from_numpy((data[i + 1:i + 1 + block_size]).astype(np.int64)) for i in ix])
    x, y = x.pin_memory().to(device, non_blocking=True), y.pin_memory().to(device, non_blocking=True)
    return x, y


iter_num = 0
best_val_loss = 1e9

meta_path = os.path.join(data_dir, 'meta.pkl')
meta_vocab_size = None
if os.path.exists(meta_path):
    with open(meta_path, 'rb') as f:
        meta_vocab_size = pickle.load(f)['vocab_size']
    if master_process:
        print(f"found vocab_size = {meta_vocab_size} (inside {meta_path})")

model_args = dict(n_layer=n_layer, n_head=n_head, n_embd=n_embd, block_size=block_size,
                  bias=bias, vocab_size=None, dropout=dropout)
checkpoint = None
if init_from == 'scratch':
    if master_process:
        print("Initializing a new model from scratch")
        if meta_vocab_size is None:
            print("defaulting to vocab_size of GPT-2 to 50304 (50257 rounded up for efficiency)")
    model_args['vocab_size'] = meta_vocab_size if meta_vocab_size is not None else 50304
    model = GPT(GPTConfig(**model_args), verbose=master_process)
elif init_from == 'resume':
    print(f"Resuming training from {out_dir}")
    checkpoint = torch.load(os.path.join(out_dir, 'ckpt.pt'), map_location='cpu')
    for k in ['n_layer', 'n_head', 'n_embd', 'block_size', 'bias', 'vocab_size']:
        model_args[k] = checkpoint['model_args'][k]
    model = GPT(GPTConfig(**model_args), verbose=master_process)
    state_dict = {k[len('_orig_mod.'):] if k.startswith('_orig_mod.') else k: v
                  for k, v in checkpoint['model'].items()}
    model.load_state_dict(state_dict)
    iter_num = checkpoint['iter_num']
    best_val_loss = checkpoint['best_val_loss']
elif init_from.startswith('gpt2'):
    print(f"Initializing from OpenAI GPT-2 weights: {init_from}")
    model = GPT.from_pretrained(init_from, dict(dropout=dropout))  # needs the HF weights (cache or network)
    for k in ['n_layer', 'n_head', 'n_embd', 'block_size', 'bias', 'vocab_size']:
        model_args[k] = getattr(model.config, k)
else:
    raise SystemExit(f"init_from={init_from!r}: expected 'scratch', 'resume' or 'gpt2*'")
if block_size < model.config.block_size:
    model.crop_block_size(block_size)
    model_args['block_size'] = block_size
model.to(device)

optimizer = model.configure_optimizers(weight_decay, learning_rate, (beta1, beta2), device_type,
                                       verbose=master_process)
if checkpoint is not None:
    optimizer.load_state_dict(checkpoint['optimizer'])
checkpoint = None

tb_writer = None
if tensorboard_log and master_process:
    from torch.utils.tensorboard import SummaryWriter
    tb_writer = SummaryWriter(log_dir=tensorboard_dir)

raw_model = model
if ddp:
    model = DDP(model, device_ids=[ddp_local_rank], bucket_dtype=bucket_dtype)


@torch.no_grad()
def estimate_loss():
    out = {}
    raw_model.eval()
    for split in ['train', 'val']:
        losses = torch.zeros(eval_iters)
        for k in range(eval_iters):
            X, Y = get_batch(split)
            with ctx:
                _, loss = raw_model(X, Y)
            losses[k] = loss.item()
        out[split] = losses.mean()
    raw_model.train()
    return out


def get_lr(it):
    if it < warmup_iters:
        return learning_rate * (it + 1) / (warmup_iters + 1)
    if it > lr_decay_iters:
        return min_lr
    decay_ratio = (it - warmup_iters) / (lr_decay_iters - warmup_iters)
    coeff = 0.5 * (1.0 + math.cos(math.pi * decay_ratio))
    return min_lr + coeff * (learning_rate - min_lr)


X, Y = get_batch('train')
t0 = time.time()
local_iter_num = 0
running_mfu = -1.0
while True:
    lr = get_lr(iter_num) if decay_lr else learning_rate
    for param_group in optimizer.param_groups:
        param_group['lr'] = lr

    if iter_num % eval_interval == 0 and master_process:
        losses = estimate_loss()
        print(f"step {iter_num}: train loss {losses['train']:.4f}, val loss {losses['val']:.4f}")
        if tb_writer is not None:
            tb_writer.add_scalar("val/loss", float(losses['val']), iter_num)
            tb_writer.add_scalar("train/eval_loss", float(losses['train']), iter_num)
        if losses['val'] < best_val_loss or always_save_checkpoint:
            best_val_loss = losses['val']
            if iter_num > 0:
                ckpt = {'model': raw_model.state_dict(), 'optimizer': optimizer.state_dict(),
                        'model_args': model_args, 'iter_num': iter_num,
                        'best_val_loss': best_val_loss, 'config': config}
                print(f"saving checkpoint to {out_dir}")
                torch.save(ckpt, os.path.join(out_dir, 'ckpt.pt'))
    if iter_num == 0 and eval_only:
        break

    for micro_step in range(gradient_accumulation_steps):
        if ddp:
            model.require_backward_grad_sync = (micro_step == gradient_accumulation_steps - 1)
        with ctx:
            logits, loss = model(X, Y)
            loss = loss / gradient_accumulation_steps
        X, Y = get_batch('train')  # prefetch the next batch while the GPU works
        loss.backward()
    if grad_clip != 0.0:
        optimizer.clip_grad_norm_(grad_clip)
    optimizer.step()
    optimizer.zero_grad(set_to_none=True)

    t1 = time.time()
    dt = t1 - t0
    t0 = t1
    if iter_num % log_interval == 0 and master_process:
        lossf = loss.item() * gradient_accumulation_steps  # device->host sync point
        if local_iter_num >= 5:
            mfu = raw_model.estimate_mfu(batch_size * gradient_accumulation_steps, dt)
            running_mfu = mfu if running_mfu == -1.0 else 0.9 * running_mfu + 0.1 * mfu
        mfu_b200 = running_mfu * 312e12 * ddp_world_size / (peak_flops * ddp_world_size) if running_mfu > 0 else -1.0
        print(f"iter {iter_num}: loss {lossf:.4f}, time {dt*1000:.2f}ms, mfu {running_mfu*100:.2f}%"
              f" | {tokens_per_iter/dt:,.0f} tok/s, mfu_b200 {mfu_b200*100:.2f}%")
        if tb_writer is not None:
            tb_writer.add_scalar("train/loss", lossf, iter_num)
            tb_writer.add_scalar("train/lr", lr, iter_num)
            tb_writer.add_scalar("perf/tokens_per_sec", tokens_per_iter / dt, iter_num)
    iter_num += 1
    local_iter_num += 1
    if iter_num > max_iters:
        break

if tb_writer is not None:
    tb_writer.close()
if ddp:
    destroy_process_group()
