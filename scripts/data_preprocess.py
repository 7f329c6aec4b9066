"""Dataset -> flat uint16 token files, as reference ``scripts/data_preprocess.py:1-63``:
0.05% validation split (seed 42), per-document ``encode_ordinary`` + EOT, sharded concatenation
into ``np.memmap(uint16)`` at ``train_path`` / ``val_path``.

Additions for offline boxes: ``--text_files a.txt b.txt`` reads local text instead of a HF dataset,
and ``--tokenizer bytes`` is a dependency-free byte-level tokenizer (ids 0..255, EOT 256) used when
tiktoken's BPE table cannot be fetched.
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from config.config import default_config as config  # noqa: E402


class ByteTokenizer:
    eot_token = 256

    def encode_ordinary(self, text: str):
        return list(text.encode("utf-8"))


def get_tokenizer(name: str):
    if name == "bytes":
        return ByteTokenizer()
    import tiktoken
    return tiktoken.get_encoding(name)


def process_and_encode(example, enc):
    ids = enc.encode_ordinary(example["text"])
    ids.append(enc.eot_token)
    return {"ids": ids, "len": len(ids)}


def write_tokens(docs, path: str, num_shards: int = 1024) -> int:
    """Concatenate tokenised documents into one uint16 memmap, shard by shard."""
    total = int(sum(d["len"] for d in docs))
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    arr = np.memmap(path, dtype=np.uint16, mode="w+", shape=(total,))
    n = len(docs)
    shards = max(1, min(num_shards, n))
    idx = 0
    for s in range(shards):
        lo, hi = s * n // shards, (s + 1) * n // shards
        if lo == hi:
            continue
        batch = np.concatenate([np.asarray(docs[i]["ids"], dtype=np.uint16) for i in range(lo, hi)])
        arr[idx:idx + len(batch)] = batch
        idx += len(batch)
    arr.flush()
    return total


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--text_files", nargs="*", default=None)
    ap.add_argument("--tokenizer", default=config.get("tokenizer_name", "gpt2"))
    ap.add_argument("--val_fraction", type=float, default=0.0005)
    args = ap.parse_args(argv)
    enc = get_tokenizer(args.tokenizer)
    if args.text_files:
        texts = []
        for f in args.text_files:
            with open(f, "r", encoding="utf-8") as fh:
                texts += [t for t in fh.read().split("\n\n") if t.strip()]
        rng = np.random.default_rng(42)
        order = rng.permutation(len(texts))
        n_val = max(1, int(round(len(texts) * args.val_fraction)))
        splits = {"val": [texts[i] for i in order[:n_val]], "train": [texts[i] for i in order[n_val:]]}
        tokenised = {k: [process_and_encode({"text": t}, enc) for t in v] for k, v in splits.items()}
    else:
        from datasets import load_dataset
        ds = load_dataset(config.get("dataset_name"))
        split = ds["train"].train_test_split(test_size=args.val_fraction, seed=42, shuffle=True)
        split["val"] = split.pop("test")
        tok = split.map(lambda ex: process_and_encode(ex, enc), remove_columns=["text"],
                        desc="tokenizing", num_proc=os.cpu_count())
        tokenised = {k: tok[k] for k in ("train", "val")}
    for name, docs in tokenised.items():
        path = config["train_path"] if name == "train" else config["val_path"]
        n = write_tokens(docs, path)
        print(f"{name}: {n:,} tokens -> {path}")


if __name__ == "__main__":
    main()
