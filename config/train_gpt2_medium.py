# GPT-2 medium 350M (24 layers, 16 heads, 1024 wide), the 8-Pod StatefulSet configuration
wandb_log = False
batch_size = 12
block_size = 1024
gradient_accumulation_steps = 40
n_layer = 24
n_head = 16
n_embd = 1024
max_iters = 600000
lr_decay_iters = 600000
eval_interval = 1000
eval_iters = 200
log_interval = 10
