"""Dataset, collate function and loader (reference dataset.py:12-68) plus a synthetic-corpus generator that emits
the same JSON schema the reference's pre_tokenize.py produces (there is no network on the build / GPU boxes).

JSON schema: {"train": [[ids...], ...], "validation": [[...]], "special_ids": {"<BOS>","<EOS>","<UNK>"},
"vocab_size": V}.  Batch layout: input_ids = [BOS, tok..., EOS-pad], target_ids = [tok..., EOS, IGNORE-pad],
position_ids = arange(max_len + 1) per row.  Every TP rank must see identical batches (same seed, no sharding).
"""
from __future__ import annotations

import json
import os
from functools import partial
from typing import List

import torch
from torch.utils.data import DataLoader, Dataset

BOS_TOKEN, EOS_TOKEN, UNK_TOKEN = "<BOS>", "<EOS>", "<UNK>"


class ShakespeareDataset(Dataset):
    def __init__(self, data_path: str, split: str, maxlen):
        assert split in ["train", "validation"], f"Expectected split to be 'train' or 'validation', but got {split}"
        assert os.path.exists(data_path), f"data file {data_path} not found"
        with open(data_path, "r") as f:
            self.data = json.load(f)
        if split not in self.data:
            raise ValueError(f"Split {split} not found in data file {data_path}. Available splits: {self.data.keys()}")
        self.maxlen = maxlen
        self.split = split
        self.bos = self.data["special_ids"][BOS_TOKEN]
        self.eos = self.data["special_ids"][EOS_TOKEN]
        self.unk = self.data["special_ids"][UNK_TOKEN]
        self.vocab_size = self.data["vocab_size"]
        self._warned = False

    def __len__(self):
        return len(self.data[self.split])

    def __getitem__(self, idx) -> List[int]:
        tokens = self.data[self.split][idx]
        if len(tokens) > self.maxlen - 1:      # reserve one position for BOS / EOS
            if not self._warned:
                print(f"Warning: sequence is longer than maxlen {self.maxlen - 1}: {len(tokens)}. Truncating...")
                self._warned = True
            tokens = tokens[:self.maxlen - 1]
        return tokens


def collate_fn(batch: List[List[int]], bos: int, eos: int, ignore_idx: int):
    max_len = max(len(x) for x in batch)
    input_ids = torch.full((len(batch), max_len + 1), fill_value=eos, dtype=torch.long)
    target_ids = torch.full((len(batch), max_len + 1), fill_value=ignore_idx, dtype=torch.long)
    for i, b in enumerate(batch):
        input_ids[i, 0] = bos
        input_ids[i, 1: len(b) + 1] = torch.tensor(b, dtype=torch.long)
        target_ids[i, :len(b)] = torch.tensor(b, dtype=torch.long)
        target_ids[i, len(b)] = eos
    position_ids = torch.arange(max_len + 1).unsqueeze(0).repeat(len(batch), 1)
    return {"input_ids": input_ids, "target_ids": target_ids, "position_ids": position_ids}


def get_dataloader(data_path: str, batch_size: int, ignore_idx: int, split: str, maxlen: int, shuffle: bool = True):
    dataset = ShakespeareDataset(data_path, split, maxlen=maxlen)
    collate = partial(collate_fn, bos=dataset.bos, eos=dataset.eos, ignore_idx=ignore_idx)
    return DataLoader(dataset, batch_size=batch_size, shuffle=shuffle, collate_fn=collate, num_workers=0,
                      pin_memory=torch.cuda.is_available())


def write_synthetic_corpus(path: str, vocab_size: int = 1024, n_train: int = 256, n_val: int = 16,
                           min_len: int = 32, max_len: int = 999, seed: int = 0, fixed_len: int = 0) -> str:
    """Token-id corpus with the reference JSON schema.  Ids are drawn from [3, vocab) with a Zipf-like skew and a
    first-order structure (so the loss can actually fall); 0/1/2 are BOS/EOS/UNK like the shipped tokenizer."""
    g = torch.Generator().manual_seed(seed)
    ranks = torch.arange(1, vocab_size - 2, dtype=torch.float64)
    probs = (1.0 / ranks) / (1.0 / ranks).sum()

    def doc():
        n = fixed_len if fixed_len > 0 else int(torch.randint(min_len, max_len + 1, (1,), generator=g))
        base = torch.multinomial(probs, n, replacement=True, generator=g) + 3
        # make every other token a deterministic function of its predecessor
        nxt = (base[:-1] * 7 + 11) % (vocab_size - 3) + 3
        base[1::2] = nxt[0::2][: base[1::2].numel()]
        return base.tolist()

    data = {"train": [doc() for _ in range(n_train)], "validation": [doc() for _ in range(n_val)],
            "special_ids": {BOS_TOKEN: 0, EOS_TOKEN: 1, UNK_TOKEN: 2}, "vocab_size": vocab_size}
    os.makedirs(os.path.dirname(os.path.abspath(path)) or "./", exist_ok=True)
    with open(path, "w") as f:
        json.dump(data, f)
    return path
