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
    """One split of a pre-tokenised corpus, held as ONE flat int64 tensor plus document offsets (CSR layout) rather
    than a list of Python lists: a 1-D slice per item, no per-token Python objects resident.

    API parity with the reference (dataset.py:12-37): constructor signature, ``bos`` / ``eos`` / ``unk`` /
    ``vocab_size`` attributes, ``__getitem__`` -> list of token ids clipped to ``maxlen - 1`` (one slot is kept for
    BOS on the input side / EOS on the target side)."""

    SPLITS = ("train", "validation")

    def __init__(self, data_path: str, split: str, maxlen):
        if split not in self.SPLITS:
            raise AssertionError(f"split must be one of {self.SPLITS}, got {split!r}")
        if not os.path.isfile(data_path):
            raise AssertionError(f"token file {data_path} does not exist (run pre_tokenize.py or "
                                 f"write_synthetic_corpus first)")
        with open(data_path, "r") as fh:
            blob = json.load(fh)
        if split not in blob:
            have = [k for k in blob if k in self.SPLITS]
            raise ValueError(f"{data_path} has no {split!r} split (it has {have})")
        docs = blob[split]
        lens = torch.tensor([len(d) for d in docs], dtype=torch.long)
        self._offsets = torch.zeros(len(docs) + 1, dtype=torch.long)
        torch.cumsum(lens, 0, out=self._offsets[1:])
        self._tokens = torch.tensor([t for d in docs for t in d], dtype=torch.long)
        ids = blob["special_ids"]
        self.bos, self.eos, self.unk = ids[BOS_TOKEN], ids[EOS_TOKEN], ids[UNK_TOKEN]
        self.vocab_size = blob["vocab_size"]
        self.maxlen, self.split = maxlen, split
        self._clip_reported = False

    def __len__(self):
        return self._offsets.numel() - 1

    def __getitem__(self, idx) -> List[int]:
        lo, hi = int(self._offsets[idx]), int(self._offsets[idx + 1])
        room = self.maxlen - 1
        if hi - lo > room:
            if not self._clip_reported:      # once per dataset, not once per sample
                print(f"note: document {idx} has {hi - lo} tokens, more than the {room} that fit maxlen="
                      f"{self.maxlen}; long documents are clipped")
                self._clip_reported = True
            hi = lo + room
        return self._tokens[lo:hi].tolist()


def collate_fn(batch: List[List[int]], bos: int, eos: int, ignore_idx: int):
    """Ragged documents -> the three (B, L + 1) int64 tensors of a training batch, L = longest document:

        input_ids    BOS d0 d1 ... d(k-1) EOS EOS ...      (EOS is the pad value on the input side)
        target_ids   d0  d1 ... d(k-1) EOS IGN IGN ...     (IGN = ignore_idx, skipped by the loss)
        position_ids 0 1 2 ... L

    Same layout as the reference's collate (dataset.py:40-55), built with one flat scatter per tensor instead of a
    Python loop of per-row tensor constructions."""
    B = len(batch)
    lens = torch.tensor([len(doc) for doc in batch], dtype=torch.long)
    width = int(lens.max()) + 1 if B else 1
    input_ids = torch.full((B, width), eos, dtype=torch.long)
    target_ids = torch.full((B, width), ignore_idx, dtype=torch.long)
    flat = torch.tensor([t for doc in batch for t in doc], dtype=torch.long)
    rows = torch.repeat_interleave(torch.arange(B), lens)
    starts = torch.cumsum(lens, 0) - lens
    cols = torch.arange(flat.numel()) - torch.repeat_interleave(starts, lens)
    input_ids[:, 0] = bos
    input_ids[rows, cols + 1] = flat
    target_ids[rows, cols] = flat
    target_ids[torch.arange(B), lens] = eos
    position_ids = torch.arange(width).expand(B, width).contiguous()
    return {"input_ids": input_ids, "target_ids": target_ids, "position_ids": position_ids}


def get_dataloader(data_path: str, batch_size: int, ignore_idx: int, split: str, maxlen: int, shuffle: bool = True):
    """Every TP rank builds this loader with the same seed and NO sampler sharding, so all ranks see identical
    batches (reference dataset.py:58-68)."""
    ds = ShakespeareDataset(data_path, split, maxlen=maxlen)
    return DataLoader(ds, batch_size=batch_size, shuffle=shuffle, num_workers=0,
                      pin_memory=torch.cuda.is_available(),
                      collate_fn=partial(collate_fn, bos=ds.bos, eos=ds.eos, ignore_idx=ignore_idx))


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
