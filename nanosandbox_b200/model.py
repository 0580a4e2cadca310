"""Drop-in for nanoGPT's ``model.py`` (GPTConfig, GPT) on hand-written sm_100a kernels.

Same public surface as the module the reference's jobs import through ``train.py``
(/root/reference/notebooks/colab_nanoGPT_companion.ipynb:70-79, 107-116):
``GPTConfig(block_size, vocab_size, n_layer, n_head, n_embd, dropout, bias)``,
``GPT(config).forward(idx, targets) -> (logits, loss)``, ``configure_optimizers``,
``get_num_params``, ``estimate_mfu``, ``crop_block_size``, ``generate``; identical module tree,
``state_dict()`` keys, weight tying and init RNG order.  The arithmetic is different: one forward
+ one manual backward over flat arenas, every op a kernel from libnanosandbox_b200.so (no torch
ops, no autograd graph inside, no CPU path).

Precision policy follows torch autocast (ATen/autocast_mode.h:832-921) without using it: inside
``torch.autocast('cuda', bfloat16)`` GEMMs and attention run on bf16 operands with fp32
accumulation, LayerNorm statistics, the residual stream, softmax-cross-entropy, gradients of
parameters and optimizer state stay fp32.  Outside autocast everything is exact fp32 (the 1e-5
parity mode).
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass

import torch
import torch.nn as nn

from . import ops
from ._lib import EPI_ACCUM, EPI_DGELU, EPI_GELU, EPI_RESIDUAL, EPI_STORE, NsbError
from .arena import ParamArena, bucket_ranges, ready_order


MAX_N_EMBD = 1024   # widest row the LayerNorm kernels take (csrc/elementwise.cu: nsb_layernorm_fwd / _bwd)


@dataclass
class GPTConfig:
    block_size: int = 1024
    vocab_size: int = 50304
    n_layer: int = 12
    n_head: int = 12
    n_embd: int = 768
    dropout: float = 0.0
    bias: bool = True


def _container_only(self, *a, **k):
    raise NsbError("sub-modules of nanosandbox_b200.GPT only hold parameters; call GPT.forward "
                   "(the fused engine drives all layers)")


class LayerNorm(nn.Module):
    def __init__(self, ndim, bias):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(ndim))
        self.bias = nn.Parameter(torch.zeros(ndim)) if bias else None

    forward = _container_only


class CausalSelfAttention(nn.Module):
    def __init__(self, config):
        super().__init__()
        assert config.n_embd % config.n_head == 0
        self.c_attn = nn.Linear(config.n_embd, 3 * config.n_embd, bias=config.bias)
        self.c_proj = nn.Linear(config.n_embd, config.n_embd, bias=config.bias)
        self.attn_dropout = nn.Dropout(config.dropout)
        self.resid_dropout = nn.Dropout(config.dropout)
        self.n_head, self.n_embd, self.dropout = config.n_head, config.n_embd, config.dropout

    forward = _container_only


class MLP(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.c_fc = nn.Linear(config.n_embd, 4 * config.n_embd, bias=config.bias)
        self.gelu = nn.GELU()
        self.c_proj = nn.Linear(4 * config.n_embd, config.n_embd, bias=config.bias)
        self.dropout = nn.Dropout(config.dropout)

    forward = _container_only


class Block(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.ln_1 = LayerNorm(config.n_embd, bias=config.bias)
        self.attn = CausalSelfAttention(config)
        self.ln_2 = LayerNorm(config.n_embd, bias=config.bias)
        self.mlp = MLP(config)

    forward = _container_only


class _Acts:
    """Activation + scratch buffers for one (B, T, precision) shape; allocated once and reused
    every micro-step (torch owns the memory, nothing is allocated inside the step)."""

    def __init__(self, cfg: GPTConfig, B: int, T: int, lowp: bool, device, train: bool = True):
        C, L, H = cfg.n_embd, cfg.n_layer, cfg.n_head
        M = B * T
        adt = torch.bfloat16 if lowp else torch.float32
        f32 = torch.float32
        new = lambda *shape, dtype=adt: torch.empty(*shape, dtype=dtype, device=device)
        self.B, self.T, self.M, self.lowp, self.adt = B, T, M, lowp, adt
        self.train = train
        self.gen = 0            # bumped by every forward that overwrites these buffers
        self.drop = None        # (p, seed, first stream id) of the forward that filled these buffers, or None
        self.proj_tmp = None    # c_proj output before dropout + residual add (allocated on first use)
        self.dmask = None       # masked copy of the residual gradient (operand of the c_proj gradient GEMMs)
        self.attn_mask = None
        self._cfg_c, self._device = C, device
        self.x = [new(M, C, dtype=f32) for _ in range(L + 1)]       # residual stream, block inputs
        self.xmid = [new(M, C, dtype=f32) for _ in range(L)]         # after attention
        self.ln1 = [new(M, C) for _ in range(L)]
        self.ln2 = [new(M, C) for _ in range(L)]
        self.stat = [[new(M, dtype=f32) for _ in range(4)] for _ in range(L)]  # mean1 rstd1 mean2 rstd2
        self.qkv = [new(M, 3 * C) for _ in range(L)]
        self.att = [new(M, C) for _ in range(L)]
        self.lse = [new(B, H, T, dtype=f32) for _ in range(L)]
        self.fch = [new(M, 4 * C) for _ in range(L)]                 # c_fc pre-activation
        self.fca = [new(M, 4 * C) for _ in range(L)]                 # gelu(c_fc)
        self.lnf = new(M, C)
        self.statf = [new(M, dtype=f32) for _ in range(2)]
        self.ldv = (cfg.vocab_size + 7) // 8 * 8                     # 16-byte rows for TMA / 128-bit IO
        self.logits_last = new(B, self.ldv)
        self.idx = None
        self.targets = None
        self.stats = None
        if not train:
            # forward(idx, targets=None): last-position logits only; no [M, V] logits, no backward scratch
            return
        self.logits = new(M, self.ldv)
        if self.ldv != cfg.vocab_size:
            self.logits.zero_()
        self.ce_lse = new(M, dtype=f32)
        self.ce_row = new(M, dtype=f32)
        # backward scratch
        self.dres = [new(M, C, dtype=f32) for _ in range(2)]
        self.dres_lowp = new(M, C) if lowp else None
        self.dh = new(M, 4 * C)
        self.dln = new(M, C)
        self.datt = new(M, C)
        self.dqkv = new(M, 3 * C)
        self.delta = new(B, H, T, dtype=f32)
        # fp32 dQ accumulator of the tensor-core attention backward (kept all-zero between calls)
        self.dq_acc = torch.zeros(M, C, dtype=f32, device=device) if (lowp and C // H == 64) else None
        self.ln_scratch = torch.empty(ops.layernorm_bwd_scratch_bytes(M, C) // 4, dtype=f32, device=device)

    def dropout_buffers(self, n_head):
        if self.proj_tmp is None:
            self.proj_tmp = torch.empty(self.M, self._cfg_c, dtype=self.adt, device=self._device)
            self.dmask = torch.empty(self.M, self._cfg_c, dtype=self.adt, device=self._device)
            # keep bytes of one layer's attention probabilities [B, H, T, T]; rewritten before every use (forward
            # and backward regenerate them), bf16 / head dim 64 only -- the exact kernels need no scratch
            self.attn_mask = (torch.empty(self.B * n_head * self.T * self.T, dtype=torch.uint8, device=self._device)
                              if (self.lowp and self._cfg_c // n_head == 64) else None)


class _FusedStep(torch.autograd.Function):
    """One autograd node for the whole model: forward = engine forward, backward = the manual
    backward that accumulates straight into the flat gradient arena (so autograd never touches
    75 separate .grad tensors)."""

    @staticmethod
    def forward(ctx, anchor, model, idx, targets):
        logits, loss = model._engine_forward(idx, targets, for_backward=True)
        ctx.model = model
        ctx.acts = model._last_acts
        ctx.gen = ctx.acts.gen
        model._step_logits = logits        # handed to GPT.forward outside the autograd node
        ctx.set_materialize_grads(False)
        return loss

    @staticmethod
    def backward(ctx, gloss):
        if gloss is None:
            return None, None, None, None
        a = ctx.acts
        if a.gen != ctx.gen:
            raise NsbError(
                "backward() of a forward whose activations have been overwritten: another forward with the "
                "same (batch, block, precision) ran in between (the engine keeps one set of activation "
                "buffers per shape). Call loss.backward() before the next forward of that shape.")
        ctx.model._engine_backward(gloss, a)
        return None, None, None, None


class GPT(nn.Module):
    def __init__(self, config: GPTConfig, verbose: bool = True):
        super().__init__()
        assert config.vocab_size is not None
        assert config.block_size is not None
        if config.n_embd > MAX_N_EMBD or config.n_embd % 4 != 0:
            raise NsbError(f"n_embd={config.n_embd}: the LayerNorm kernels need a multiple of 4 up to {MAX_N_EMBD}")
        self.config = config
        # same constructors in the same order as nanoGPT -> same consumption of the global RNG,
        # so get_batch() draws the same indices afterwards (hard parts)
        self.transformer = nn.ModuleDict(dict(
            wte=nn.Embedding(config.vocab_size, config.n_embd),
            wpe=nn.Embedding(config.block_size, config.n_embd),
            drop=nn.Dropout(config.dropout),
            h=nn.ModuleList([Block(config) for _ in range(config.n_layer)]),
            ln_f=LayerNorm(config.n_embd, bias=config.bias),
        ))
        self.lm_head = nn.Linear(config.n_embd, config.vocab_size, bias=False)
        self.transformer.wte.weight = self.lm_head.weight
        self.apply(self._init_weights)
        for pn, p in self.named_parameters():
            if pn.endswith("c_proj.weight"):
                torch.nn.init.normal_(p, mean=0.0, std=0.02 / math.sqrt(2 * config.n_layer))
        if verbose:
            print("number of parameters: %.2fM" % (self.get_num_params() / 1e6,))
        self._arena: ParamArena | None = None
        self._acts = {}
        self._infer_acts = None
        self._dropout_calls = 0
        self._step_logits = None
        self._grad_sync = None          # set by nanosandbox_b200.ddp.DistributedDataParallel
        # bf16 path: GELU as its own memory-bound kernels (faster on B200 than the fused GEMM epilogues,
        # whose eight epilogue warps cannot hide the erf math behind a K = 768 mainloop)
        self.split_gelu = os.environ.get("NSB_SPLIT_GELU", "1") != "0"
        # NSB_SPLIT_GELU=fwd: forward through the fused epilogue (c_fc GEMM writes h and gelu(h)), backward split
        self.fuse_gelu_fwd = os.environ.get("NSB_SPLIT_GELU", "1") == "fwd"
        # Training forward (bf16): the loss kernel also leaves softmax - onehot in the logits buffer (one
        # pass over the 1.2 GB of logits instead of two).  The logits no longer exist after that, so such
        # a forward returns (None, loss).  Set to False if you read the logits of a training forward.
        self.fuse_loss_backward = True
        # Weight-gradient GEMMs run on a side stream: nothing in the backward chain depends on them, so
        # their CTAs fill the SMs that the chain's kernels leave idle in their ramp-up and tail.
        self.overlap_wgrad = os.environ.get("NSB_OVERLAP_WGRAD", "1") != "0"
        self._wgrad_stream = None
        self._micro_step_has_grads = False

    # ------------------------------------------------------------------ nanoGPT surface
    def get_num_params(self, non_embedding=True):
        n = sum(p.numel() for p in self.parameters())
        if non_embedding:
            n -= self.transformer.wpe.weight.numel()
        return n

    @staticmethod
    def _init_weights(module):
        if isinstance(module, nn.Linear):
            torch.nn.init.normal_(module.weight, mean=0.0, std=0.02)
            if module.bias is not None:
                torch.nn.init.zeros_(module.bias)
        elif isinstance(module, nn.Embedding):
            torch.nn.init.normal_(module.weight, mean=0.0, std=0.02)

    def crop_block_size(self, block_size):
        assert block_size <= self.config.block_size
        self.config.block_size = block_size
        self._arena = None
        self._acts = {}
        self._infer_acts = None
        self.transformer.wpe.weight = nn.Parameter(self.transformer.wpe.weight.detach()[:block_size].clone())

    @classmethod
    def from_hf_state_dict(cls, sd_hf, config_args, verbose=False):
        """Builds a GPT from a Hugging Face GPT-2 state dict: same key names as nanoGPT; the four Conv1D
        weights (attn.c_attn, attn.c_proj, mlp.c_fc, mlp.c_proj) are stored transposed there."""
        model = cls(GPTConfig(**config_args), verbose=verbose)
        sd = model.state_dict()
        transposed = ("attn.c_attn.weight", "attn.c_proj.weight", "mlp.c_fc.weight", "mlp.c_proj.weight")
        keys_hf = [k for k in sd_hf if not k.endswith((".attn.masked_bias", ".attn.bias"))]
        keys = [k for k in sd if not k.endswith(".attn.bias")]
        assert len(keys_hf) == len(keys), f"mismatched keys: {len(keys_hf)} != {len(keys)}"
        with torch.no_grad():
            for k in keys_hf:
                src = sd_hf[k].t() if k.endswith(transposed) else sd_hf[k]
                assert src.shape == sd[k].shape, (k, tuple(src.shape), tuple(sd[k].shape))
                sd[k].copy_(src)
        return model

    @classmethod
    def from_pretrained(cls, model_type, override_args=None):
        """nanoGPT's from_pretrained: loads OpenAI GPT-2 weights through the Hugging Face hub (needs the
        weights in the local HF cache or network access)."""
        assert model_type in {"gpt2", "gpt2-medium", "gpt2-large", "gpt2-xl"}
        if model_type in ("gpt2-large", "gpt2-xl"):
            raise NsbError(f"from_pretrained({model_type!r}): n_embd > {MAX_N_EMBD} is not supported by the "
                           "LayerNorm kernels of this library (gpt2 and gpt2-medium are)")
        override_args = override_args or {}
        assert all(k == "dropout" for k in override_args)
        from transformers import GPT2LMHeadModel
        config_args = {"gpt2": dict(n_layer=12, n_head=12, n_embd=768),
                       "gpt2-medium": dict(n_layer=24, n_head=16, n_embd=1024),
                       "gpt2-large": dict(n_layer=36, n_head=20, n_embd=1280),
                       "gpt2-xl": dict(n_layer=48, n_head=25, n_embd=1600)}[model_type]
        print("forcing vocab_size=50257, block_size=1024, bias=True")
        config_args.update(vocab_size=50257, block_size=1024, bias=True)
        if "dropout" in override_args:
            config_args["dropout"] = override_args["dropout"]
        try:
            hf = GPT2LMHeadModel.from_pretrained(model_type)
        except Exception as exc:
            raise NsbError(f"from_pretrained({model_type!r}): the Hugging Face weights are not available "
                           f"offline ({type(exc).__name__})") from exc
        return cls.from_hf_state_dict(hf.state_dict(), config_args)

    def configure_optimizers(self, weight_decay, learning_rate, betas, device_type, verbose=True):
        from .optim import FusedAdamW
        if device_type != "cuda":
            raise NsbError("configure_optimizers: device_type must be 'cuda' (no CPU path)")
        arena = self.arena()
        params = {n: p for n, p in self.named_parameters() if p.requires_grad}
        decay = [p for p in params.values() if p.dim() >= 2]
        nodecay = [p for p in params.values() if p.dim() < 2]
        if verbose:
            print(f"num decayed parameter tensors: {len(decay)}, with {sum(p.numel() for p in decay):,} parameters")
            print(f"num non-decayed parameter tensors: {len(nodecay)}, with {sum(p.numel() for p in nodecay):,} parameters")
            print("using fused AdamW: True")
        groups = [{"params": decay, "weight_decay": weight_decay}, {"params": nodecay, "weight_decay": 0.0}]
        return FusedAdamW(groups, arena, lr=learning_rate, betas=betas)

    def estimate_mfu(self, fwdbwd_per_iter, dt, peak_flops: float = 312e12):
        """nanoGPT's formula (6N + 12LHQT per token).  peak_flops defaults to nanoGPT's A100
        constant so logs stay comparable; train.py also prints MFU against the B200 peak."""
        cfg = self.config
        N = self.get_num_params()
        L, H, Q, T = cfg.n_layer, cfg.n_head, cfg.n_embd // cfg.n_head, cfg.block_size
        flops_per_iter = (6 * N + 12 * L * H * Q * T) * T * fwdbwd_per_iter
        return flops_per_iter / dt / peak_flops

    @torch.no_grad()
    def generate(self, idx, max_new_tokens, temperature=1.0, top_k=None):
        for _ in range(max_new_tokens):
            cond = idx if idx.size(1) <= self.config.block_size else idx[:, -self.config.block_size:]
            logits, _ = self(cond)
            logits = logits[:, -1, :].float() / temperature
            if top_k is not None:
                v, _ = torch.topk(logits, min(top_k, logits.size(-1)))
                logits[logits < v[:, [-1]]] = -float("inf")
            probs = torch.softmax(logits, dim=-1)
            idx = torch.cat((idx, torch.multinomial(probs, num_samples=1)), dim=1)
        return idx

    # ------------------------------------------------------------------ arena plumbing
    def arena(self) -> ParamArena:
        """Move the parameters into the flat arena (once, after .to('cuda')); Parameter objects
        keep their identity, their storage becomes a view of the arena and .grad a view of the
        gradient arena."""
        dev = self.lm_head.weight.device
        if self._arena is not None and self._arena.device == dev:
            return self._arena
        if dev.type != "cuda":
            raise NsbError("nanosandbox_b200.GPT runs on CUDA only: call model.to('cuda') first "
                           "(there is no CPU fallback; the CPU restatement lives in oracle/)")
        named = ready_order(list(self.named_parameters()), self.config.n_layer)
        arena = ParamArena(named, dev)
        with torch.no_grad():
            for name, p in named:
                view = arena.p(name)
                view.copy_(p.detach().to(torch.float32))
                p.data = view
                p.grad = arena.g(name)
        self._names = {id(p): n for n, p in named}
        self._arena = arena
        self._acts = {}
        self._infer_acts = None
        self._has_bias = self.config.bias
        self.bucket_ranges = bucket_ranges(arena, self.config.n_layer, self.config.bias)
        return arena

    def _apply(self, fn, *a, **k):
        out = super()._apply(fn, *a, **k)
        ar = self._arena
        if ar is not None and self.lm_head.weight.data_ptr() != ar.p("transformer.wte.weight").data_ptr():
            self._arena = None  # storage moved (e.g. .to(other device)): rebuild lazily
            self._acts = {}
            self._infer_acts = None
        return out

    def load_state_dict(self, state_dict, *a, **k):
        out = super().load_state_dict(state_dict, *a, **k)
        if self._arena is not None:
            self._arena.lowp_version = -1
        return out

    def zero_grad(self, set_to_none: bool = True):
        if self._arena is not None:
            self._arena.grads.zero_()  # .grad stays a view of the arena
        else:
            super().zero_grad(set_to_none)

    def _refresh_lowp(self, arena: ParamArena, st):
        ver = arena.params._version
        if arena.lowp_version != ver:
            ops.cast_f32_to_bf16(arena.params, arena.lowp, 1.0, st)
            arena.lowp_version = ver

    # ------------------------------------------------------------------ forward
    def forward(self, idx, targets=None):
        b, t = idx.size()
        assert t <= self.config.block_size, \
            f"Cannot forward sequence of length {t}, block size is only {self.config.block_size}"
        if targets is not None and torch.is_grad_enabled() and self.lm_head.weight.requires_grad:
            loss = _FusedStep.apply(self.lm_head.weight, self, idx, targets)
            logits, self._step_logits = self._step_logits, None
            return logits, loss
        return self._engine_forward(idx, targets)

    def _mode_lowp(self) -> bool:
        if torch.is_autocast_enabled("cuda"):
            dt = torch.get_autocast_dtype("cuda")
            if dt == torch.bfloat16:
                return True
            if dt == torch.float32:
                return False
            raise NotImplementedError(f"autocast dtype {dt} is not supported (bfloat16 or float32)")
        return False

    _MAX_TRAIN_SHAPES = 4

    def _get_acts(self, B, T, lowp, train=True) -> _Acts:
        """Activation buffers for this shape.  Training/eval shapes (targets given) are cached, a few
        at a time (train batch + eval batch); the inference path (targets=None, e.g. generate(),
        which calls forward with a new T for every token) keeps ONE forward-only set and replaces it
        when the shape changes, so its memory does not grow with the number of generated tokens."""
        if not train:
            a = self._infer_acts
            if a is None or (a.B, a.T, a.lowp) != (B, T, lowp):
                self._infer_acts = None      # release before allocating the replacement
                a = self._infer_acts = _Acts(self.config, B, T, lowp, self.lm_head.weight.device, train=False)
            return a
        key = (B, T, lowp)
        a = self._acts.pop(key, None)
        if a is None:
            while len(self._acts) >= self._MAX_TRAIN_SHAPES:
                self._acts.pop(next(iter(self._acts)))      # least recently used
            a = _Acts(self.config, B, T, lowp, self.lm_head.weight.device)
        self._acts[key] = a                                 # most recently used last
        return a

    def _engine_forward(self, idx, targets, for_backward=False):
        cfg = self.config
        arena = self.arena()
        if not idx.is_cuda:
            raise NsbError("idx must be a CUDA tensor (no CPU path)")
        lowp = self._mode_lowp()
        B, T = idx.shape
        C, L, H, V = cfg.n_embd, cfg.n_layer, cfg.n_head, cfg.vocab_size
        D = C // H
        M = B * T
        a = self._get_acts(B, T, lowp, train=targets is not None)
        a.gen += 1               # a backward() of an earlier forward on these buffers is now invalid
        st = ops.stream_handle()
        if lowp:
            self._refresh_lowp(arena, st)
        idx = idx.contiguous()
        a.idx = idx
        W = lambda n: arena.w(n, lowp)
        P = arena.p
        bias = (lambda n: P(n)) if cfg.bias else (lambda n: None)
        # dropout (training mode, config.dropout > 0): the four nn.Dropout sites of nanoGPT.  Masks are Philox
        # functions of (seed, stream id, element); every site of every forward gets its own stream id, which
        # backward reuses.  The seed is the CUDA generator's (torch.manual_seed sets it, per rank in train.py).
        pdrop = float(cfg.dropout) if (self.training and cfg.dropout > 0.0) else 0.0
        if pdrop > 0.0:
            a.dropout_buffers(H)
            a.drop = (pdrop, torch.cuda.initial_seed() & ((1 << 63) - 1), self._dropout_calls * (3 * L + 1))
            self._dropout_calls += 1
        else:
            a.drop = None
        seed, sid0 = (a.drop[1], a.drop[2]) if a.drop else (0, 0)

        def proj_residual(l, x_in, k_in, wname, res, out, site):
            # out = res + dropout(x_in @ W^T + b): the residual epilogue when there is no dropout, else the
            # GEMM stores its result and the dropout kernel does the add
            if a.drop is None:
                ops.gemm(M, C, k_in, x_in, k_in, True, W(wname + ".weight"), k_in, True, out, C,
                         EPI_RESIDUAL, aux=res, bias=bias(wname + ".bias"), stream=st)
            else:
                ops.gemm(M, C, k_in, x_in, k_in, True, W(wname + ".weight"), k_in, True, a.proj_tmp, C,
                         EPI_STORE, bias=bias(wname + ".bias"), stream=st)
                ops.dropout(a.proj_tmp, out, pdrop, seed, sid0 + site + 3 * l, residual=res, stream=st)

        ops.embedding_fwd(idx, P("transformer.wte.weight"), P("transformer.wpe.weight"), a.x[0], st)
        if a.drop:
            ops.dropout(a.x[0], a.x[0], pdrop, seed, sid0, stream=st)
        for l in range(L):
            pre = f"transformer.h.{l}."
            s = a.stat[l]
            ops.layernorm_fwd(a.x[l], P(pre + "ln_1.weight"), bias(pre + "ln_1.bias"), a.ln1[l], s[0], s[1], 1e-5, st)
            ops.gemm(M, 3 * C, C, a.ln1[l], C, True, W(pre + "attn.c_attn.weight"), C, True, a.qkv[l], 3 * C,
                     EPI_STORE, bias=bias(pre + "attn.c_attn.bias"), stream=st)
            if a.drop:
                ops.attention_fwd_dropout(a.qkv[l], a.att[l], a.lse[l], B, T, H, D, pdrop, seed, sid0 + 1 + 3 * l,
                                          a.attn_mask, st)
            else:
                ops.attention_fwd(a.qkv[l], a.att[l], a.lse[l], B, T, H, D, st)
            proj_residual(l, a.att[l], C, pre + "attn.c_proj", a.x[l], a.xmid[l], 2)
            ops.layernorm_fwd(a.xmid[l], P(pre + "ln_2.weight"), bias(pre + "ln_2.bias"), a.ln2[l], s[2], s[3], 1e-5, st)
            if lowp and self.split_gelu and not self.fuse_gelu_fwd:
                ops.gemm(M, 4 * C, C, a.ln2[l], C, True, W(pre + "mlp.c_fc.weight"), C, True, a.fch[l], 4 * C,
                         EPI_STORE, bias=bias(pre + "mlp.c_fc.bias"), stream=st)
                ops.gelu_fwd(a.fch[l], a.fca[l], st)
            else:
                ops.gemm(M, 4 * C, C, a.ln2[l], C, True, W(pre + "mlp.c_fc.weight"), C, True, a.fch[l], 4 * C,
                         EPI_GELU, C2=a.fca[l], bias=bias(pre + "mlp.c_fc.bias"), stream=st)
            proj_residual(l, a.fca[l], 4 * C, pre + "mlp.c_proj", a.xmid[l], a.x[l + 1], 3)
        ops.layernorm_fwd(a.x[L], P("transformer.ln_f.weight"), bias("transformer.ln_f.bias"), a.lnf,
                          a.statf[0], a.statf[1], 1e-5, st)
        wte = W("transformer.wte.weight")
        if targets is None:
            # inference: lm_head on the last position only; rows strided by T*C, read in place
            last = a.lnf.view(B, T, C)[:, -1, :]
            ops.gemm(B, V, C, last, T * C, True, wte, C, True, a.logits_last, a.ldv, EPI_STORE, stream=st)
            return a.logits_last[:, :V].view(B, 1, V), None
        targets = targets.contiguous().view(-1)
        a.targets = targets
        ops.gemm(M, V, C, a.lnf, C, True, wte, C, True, a.logits, a.ldv, EPI_STORE, stream=st)
        a.stats = torch.empty(2, dtype=torch.float32, device=idx.device)
        a.ce_fused = bool(for_backward and self.fuse_loss_backward and ops.cross_entropy_fused_ok(a.logits))
        if a.ce_fused:
            ops.cross_entropy_fwd_bwd(a.logits, targets, a.ce_lse, a.ce_row, a.stats, V, st)
        else:
            ops.cross_entropy_fwd(a.logits, targets, a.ce_lse, a.ce_row, a.stats, V, st)
        self._last_acts = a
        if a.ce_fused:
            # the buffer now holds softmax - onehot, not logits: hand back None rather than a tensor
            # that is not what its name says (set GPT.fuse_loss_backward = False to get the logits)
            return None, a.stats[0]
        # NOTE: the logits buffer is reused: backward() turns it into dlogits in place and the next
        # forward overwrites it.  Clone it if you need the values later.
        return a.logits[:, :V].view(B, T, V), a.stats[0]

    # ------------------------------------------------------------------ backward
    def _engine_backward(self, gloss, a: _Acts):
        cfg = self.config
        arena = self._arena
        lowp = a.lowp
        B, T, M = a.B, a.T, a.M
        C, L, H, V = cfg.n_embd, cfg.n_layer, cfg.n_head, cfg.vocab_size
        D = C // H
        st = ops.stream_handle()
        gs = gloss.to(torch.float32).contiguous() if gloss.dtype != torch.float32 else gloss.contiguous()
        W = lambda n: arena.w(n, lowp)
        P, G = arena.p, arena.g
        has_bias = cfg.bias
        sync = self._grad_sync
        act = (lambda t32, tl: tl) if lowp else (lambda t32, tl: t32)
        drop = a.drop

        # the gradient all-reduce of this backward will share the GPU with it: let the GEMMs' running CTAs absorb
        # the tiles of CTAs that find their SM taken (dynamic tile counter) until the buckets are joined
        overlapped_comm = sync is not None and getattr(sync, "require_backward_grad_sync", True)
        if overlapped_comm:
            ops.gemm_dynamic_tiles(True)
        main = torch.cuda.current_stream()
        side = None
        # A backward whose buckets are reduced as it goes keeps every weight-gradient GEMM on the compute stream: on
        # 8 GPUs one replica was seen to drift from the others in a few rows of the LAST weight gradient of the first
        # bucket (h.11.attn.c_attn; bench.py ddp_check) - the signature of that GEMM's
        # accumulation landing after the bucket had been reduced and written back.  The side stream buys 0.3 % of a
        # micro-step; one micro-step per optimizer step does without it.
        if self.overlap_wgrad and not overlapped_comm:
            if self._wgrad_stream is None or self._wgrad_stream.device != main.device:
                self._wgrad_stream = torch.cuda.Stream(device=main.device)
            side = self._wgrad_stream
        st_w = side.cuda_stream if side is not None else st
        pending = {}     # buffer name -> event: the side stream has finished reading that dY buffer

        def wgrad(dy, dy_ld, x, x_ld, name, n_out, n_in, alpha=None, reads=None):
            # dW[n_out, n_in] += dy^T x ; both operands read "transposed" in place (MN-major)
            if side is not None:
                side.wait_event(main.record_event())          # dy is complete on the main stream
            ops.gemm(n_out, n_in, M, dy, dy_ld, False, x, x_ld, False, G(name + ".weight"), n_in, EPI_ACCUM,
                     alpha=alpha, accumulate=True, stream=st_w)
            if has_bias and (name + ".bias") in arena.slots:
                ops.colsum(dy[:, :n_out], G(name + ".bias"), True, st_w)
            if side is not None and reads is not None:
                pending[reads] = side.record_event()

        def before_overwrite(buf):
            # the main stream is about to rewrite a dY buffer a weight-gradient GEMM may still be reading
            ev = pending.pop(buf, None)
            if ev is not None:
                main.wait_event(ev)

        def join_side():
            if side is not None:
                main.wait_event(side.record_event())
                pending.clear()

        # ---- head: dlogits in place, tied lm_head wgrad + dgrad, ln_f
        if a.ce_fused:
            # the logits buffer already holds softmax - onehot; gscale / n_valid rides on the two GEMMs
            head_alpha = (gs.reshape(()) / a.stats[1]).reshape(1)
        else:
            ops.cross_entropy_bwd(a.logits, a.targets, a.ce_lse, a.stats, gs, V, st)
            head_alpha = None
        a.head_alpha = head_alpha   # stays alive until the side stream has used it
        wgrad(a.logits, a.ldv, a.lnf, C, "transformer.wte", V, C, alpha=head_alpha)
        ops.gemm(M, C, V, a.logits, a.ldv, True, W("transformer.wte.weight"), C, False, a.dln, C, EPI_STORE,
                 alpha=head_alpha, stream=st)
        cur = 0
        ops.layernorm_bwd(a.dln, a.x[L], P("transformer.ln_f.weight"), a.statf[0], a.statf[1], None,
                          a.dres[cur], a.dres_lowp, G("transformer.ln_f.weight"),
                          G("transformer.ln_f.bias") if has_bias else None, True, a.ln_scratch, st)
        for i, l in enumerate(range(L - 1, -1, -1)):
            pre = f"transformer.h.{l}."
            s = a.stat[l]
            dres, dres_a = a.dres[cur], act(a.dres[cur], a.dres_lowp)
            rd = "dres"
            if drop:      # gradient through mlp.dropout: the c_proj GEMMs see the masked copy
                before_overwrite("dmask")
                ops.dropout(dres, a.dmask, drop[0], drop[1], drop[2] + 3 + 3 * l, stream=st)
                dres_a, rd = a.dmask, "dmask"
            # MLP
            wgrad(dres_a, C, a.fca[l], 4 * C, pre + "mlp.c_proj", C, 4 * C, reads=rd)
            before_overwrite("dh")
            if lowp and self.split_gelu:
                ops.gemm(M, 4 * C, C, dres_a, C, True, W(pre + "mlp.c_proj.weight"), 4 * C, False, a.dh, 4 * C,
                         EPI_STORE, stream=st)
                ops.gelu_bwd(a.dh, a.fch[l], st)
            else:
                ops.gemm(M, 4 * C, C, dres_a, C, True, W(pre + "mlp.c_proj.weight"), 4 * C, False, a.dh, 4 * C,
                         EPI_DGELU, aux=a.fch[l], stream=st)
            wgrad(a.dh, 4 * C, a.ln2[l], C, pre + "mlp.c_fc", 4 * C, C, reads="dh")
            ops.gemm(M, C, 4 * C, a.dh, 4 * C, True, W(pre + "mlp.c_fc.weight"), C, False, a.dln, C, EPI_STORE, stream=st)
            before_overwrite("dres")
            ops.layernorm_bwd(a.dln, a.xmid[l], P(pre + "ln_2.weight"), s[2], s[3], dres, a.dres[cur ^ 1],
                              a.dres_lowp, G(pre + "ln_2.weight"), G(pre + "ln_2.bias") if has_bias else None,
                              True, a.ln_scratch, st)
            cur ^= 1
            dres, dres_a = a.dres[cur], act(a.dres[cur], a.dres_lowp)
            rd = "dres"
            if drop:      # gradient through attn.resid_dropout
                before_overwrite("dmask")
                ops.dropout(dres, a.dmask, drop[0], drop[1], drop[2] + 2 + 3 * l, stream=st)
                dres_a, rd = a.dmask, "dmask"
            # attention
            wgrad(dres_a, C, a.att[l], C, pre + "attn.c_proj", C, C, reads=rd)
            ops.gemm(M, C, C, dres_a, C, True, W(pre + "attn.c_proj.weight"), C, False, a.datt, C, EPI_STORE, stream=st)
            before_overwrite("dqkv")
            if drop:
                ops.attention_bwd_dropout(a.qkv[l], a.att[l], a.datt, a.lse[l], a.dqkv, a.delta, B, T, H, D,
                                          drop[0], drop[1], drop[2] + 1 + 3 * l, a.attn_mask, st)
            else:
                ops.attention_bwd(a.qkv[l], a.att[l], a.datt, a.lse[l], a.dqkv, a.delta, B, T, H, D, a.dq_acc, st)
            wgrad(a.dqkv, 3 * C, a.ln1[l], C, pre + "attn.c_attn", 3 * C, C, reads="dqkv")
            ops.gemm(M, C, 3 * C, a.dqkv, 3 * C, True, W(pre + "attn.c_attn.weight"), C, False, a.dln, C, EPI_STORE, stream=st)
            before_overwrite("dres")
            ops.layernorm_bwd(a.dln, a.x[l], P(pre + "ln_1.weight"), s[0], s[1], dres, a.dres[cur ^ 1],
                              a.dres_lowp, G(pre + "ln_1.weight"), G(pre + "ln_1.bias") if has_bias else None,
                              True, a.ln_scratch, st)
            cur ^= 1
            if sync is not None and getattr(sync, "require_backward_grad_sync", True):
                join_side()          # the bucket's weight gradients come from the side stream
            if sync is not None:
                sync.bucket_ready(i)
        join_side()                  # wte: lm_head weight gradient before the embedding scatter; also the final join
        if drop:                     # gradient through transformer.drop
            ops.dropout(a.dres[cur], a.dres[cur], drop[0], drop[1], drop[2], stream=st)
        ops.embedding_bwd(a.idx, a.dres[cur], G("transformer.wte.weight"), G("transformer.wpe.weight"), True, st)
        if sync is not None:
            sync.bucket_ready(L)
            sync.finish_backward()
        if overlapped_comm:
            ops.gemm_dynamic_tiles(False)
