"""nanoGPT-compatible import path: ``from model import GPTConfig, GPT``."""
from nanosandbox_b200.model import GPT, GPTConfig, LayerNorm, CausalSelfAttention, MLP, Block  # noqa: F401
