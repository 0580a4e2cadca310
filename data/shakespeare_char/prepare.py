"""
Character-level dataset in nanoGPT's on-disk format: train.bin / val.bin (uint16 token ids, 90 / 10 split)
and meta.pkl (vocab_size, stoi, itos) next to this file.  The reference's download Job runs nanoGPT's
prepare step before training (/root/reference/README.md:48-53; notebook :50-57).

    python data/shakespeare_char/prepare.py                 # uses ./input.txt, fetches tiny-shakespeare if absent
    python data/shakespeare_char/prepare.py path/to/any.txt

Without network access put a text file at data/shakespeare_char/input.txt (or pass a path); for throughput
runs `train.py --synthetic=True` writes a seeded synthetic stream in the same format instead.
"""
import os
import pickle
import sys

import numpy as np

here = os.path.dirname(os.path.abspath(__file__))
src = sys.argv[1] if len(sys.argv) > 1 else os.path.join(here, 'input.txt')
if not os.path.exists(src):
    url = 'https://raw.githubusercontent.com/karpathy/char-rnn/master/data/tinyshakespeare/input.txt'
    try:
        import urllib.request
        print(f"downloading {url}")
        urllib.request.urlretrieve(url, src)
    except Exception as e:  # offline box
        raise SystemExit(f"{src} not found and the download failed ({e}); place a text file there or pass a path")
with open(src, 'r', encoding='utf-8') as f:
    text = f.read()
print(f"length of dataset in characters: {len(text):,}")

chars = sorted(set(text))
vocab_size = len(chars)
print("all the unique characters:", ''.join(chars))
print(f"vocab size: {vocab_size:,}")
if vocab_size > 65535:
    raise SystemExit("more than 65535 distinct characters: does not fit the uint16 token format")
stoi = {ch: i for i, ch in enumerate(chars)}
itos = {i: ch for i, ch in enumerate(chars)}

ids = np.fromiter((stoi[c] for c in text), dtype=np.uint16, count=len(text))
n_train = int(len(ids) * 0.9)
train_ids, val_ids = ids[:n_train], ids[n_train:]
print(f"train has {len(train_ids):,} tokens")
print(f"val has {len(val_ids):,} tokens")
train_ids.tofile(os.path.join(here, 'train.bin'))
val_ids.tofile(os.path.join(here, 'val.bin'))
with open(os.path.join(here, 'meta.pkl'), 'wb') as f:
    pickle.dump({'vocab_size': vocab_size, 'itos': itos, 'stoi': stoi}, f)
