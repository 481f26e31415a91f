"""``from models.layers import ...`` (reference models/layers.py:14-155: RowParallelLinear :14-55, ColumnParallelLinear :58-100,
ParallelVocabularyEmbedding :103-141, RMSNorm :144-155; implementation in distributed_pytorch_from_scratch_b200/models/layers.py)."""
from distributed_pytorch_from_scratch_b200.models.layers import (  # noqa: F401
    ColumnParallelLinear, LayerNorm, ParallelVocabEmbedding, ParallelVocabularyEmbedding, RMSNorm,
    RowParallelLinear)
