"""``from models.layers import ...`` (reference models/layers.py)."""
from distributed_pytorch_from_scratch_b200.models.layers import (  # noqa: F401
    ColumnParallelLinear, LayerNorm, ParallelVocabEmbedding, ParallelVocabularyEmbedding, RMSNorm,
    RowParallelLinear)
