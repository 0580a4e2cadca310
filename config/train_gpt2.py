# GPT-2 124M, ~0.5M tokens per optimizer step: 12 batch * 1024 block * 40 accumulation steps
# (train.py divides the accumulation steps by WORLD_SIZE under torchrun)
wandb_log = False
wandb_project = 'owt'
wandb_run_name = 'gpt2-124M'
batch_size = 12
block_size = 1024
gradient_accumulation_steps = 40
max_iters = 600000
lr_decay_iters = 600000
eval_interval = 1000
eval_iters = 200
log_interval = 10
weight_decay = 1e-1
