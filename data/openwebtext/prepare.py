"""
GPT-2 BPE dataset in nanoGPT's on-disk format: train.bin / val.bin of uint16 token ids next to this file (no
meta.pkl: train.py then uses vocab_size 50304).  This is the step the reference's backlog asks a Job for
(/root/reference/scripts/gh_sync.ps1:139-148) before `train.py config/train_gpt2.py`.

    python data/openwebtext/prepare.py                      # the `openwebtext` dataset through Hugging Face datasets
    python data/openwebtext/prepare.py docs/*.txt           # or any local text files, one document per file
    python data/openwebtext/prepare.py --val-fraction 0.0005 --num-proc 8 ...

Every document is encoded with tiktoken's "gpt2" encoding (encode_ordinary: special tokens in the text are
treated as text) and followed by <|endoftext|> (50256), as nanoGPT does.  Both sources need files that are
not on an offline box (the dataset, tiktoken's BPE table): the script says which one is missing and stops;
for throughput runs `train.py --synthetic=True` writes a seeded synthetic stream in the same format instead.
"""
import argparse
import os
import sys

import numpy as np

here = os.path.dirname(os.path.abspath(__file__))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("files", nargs="*", help="local text files (default: the Hugging Face `openwebtext` dataset)")
    ap.add_argument("--val-fraction", type=float, default=0.0005)
    ap.add_argument("--num-proc", type=int, default=max(1, (os.cpu_count() or 2) // 2))
    ap.add_argument("--seed", type=int, default=2357)
    ap.add_argument("--out", default=here)
    args = ap.parse_args()

    try:
        import tiktoken
        enc = tiktoken.get_encoding("gpt2")
        eot = enc.eot_token
    except Exception as e:  # no network and no cached BPE table
        raise SystemExit(f"tiktoken's gpt2 encoding is not available ({type(e).__name__}: {e}); "
                         "it needs its BPE table in the tiktoken cache (TIKTOKEN_CACHE_DIR) or network access")

    def encode(text):
        ids = enc.encode_ordinary(text)
        ids.append(eot)
        return ids

    if args.files:
        rng = np.random.RandomState(args.seed)
        docs = {"train": [], "val": []}
        for path in args.files:
            with open(path, "r", encoding="utf-8") as f:
                docs["val" if rng.rand() < args.val_fraction else "train"].append(f.read())
        if not docs["val"]:                       # tiny inputs: hold the last document out
            docs["val"].append(docs["train"].pop() if len(docs["train"]) > 1 else docs["train"][0])
        for split, texts in docs.items():
            ids = np.fromiter((t for text in texts for t in encode(text)), dtype=np.uint16)
            ids.tofile(os.path.join(args.out, f"{split}.bin"))
            print(f"{split}.bin: {len(ids):,} tokens from {len(texts)} documents")
        return

    try:
        from datasets import load_dataset
    except ImportError:
        raise SystemExit("the Hugging Face `datasets` package is not installed: pass local text files instead")
    try:
        ds = load_dataset("openwebtext", num_proc=args.num_proc)
    except Exception as e:
        raise SystemExit(f"could not load the openwebtext dataset ({type(e).__name__}: {e}); pass local text files instead")
    split = ds["train"].train_test_split(test_size=args.val_fraction, seed=args.seed, shuffle=True)
    split["val"] = split.pop("test")
    tok = split.map(lambda ex: {"ids": encode(ex["text"]), "len": len(enc.encode_ordinary(ex["text"])) + 1},
                    remove_columns=["text"], desc="tokenizing", num_proc=args.num_proc)
    for name, dset in tok.items():
        n = int(np.sum(dset["len"], dtype=np.uint64))
        arr = np.memmap(os.path.join(args.out, f"{name}.bin"), dtype=np.uint16, mode="w+", shape=(n,))
        idx = 0
        shards = min(1024, max(1, len(dset)))
        for b in range(shards):
            batch = dset.shard(num_shards=shards, index=b, contiguous=True).with_format("numpy")
            chunk = np.concatenate(batch["ids"]) if len(batch) else np.empty(0, dtype=np.uint16)
            arr[idx:idx + len(chunk)] = chunk
            idx += len(chunk)
        arr.flush()
        print(f"{name}.bin: {n:,} tokens")


if __name__ == "__main__":
    sys.exit(main())
