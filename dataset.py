"""``from dataset import get_dataloader`` (reference dataset.py:12-68)."""
from distributed_pytorch_from_scratch_b200.utils.data import (  # noqa: F401
    ShakespeareDataset, collate_fn, get_dataloader, write_synthetic_corpus)
