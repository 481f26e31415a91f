"""``from models.model import Transformer`` (reference models/model.py:17-158: RoPE helpers :17-46, Attention :49-78, FFN :81-95,
DecoderLayer :98-121, Transformer :124-158; VallinaTransformer is the oracle its tests/test_transformers.py:14 expects)."""
from distributed_pytorch_from_scratch_b200.models.model import (  # noqa: F401
    FFN, Attention, DecoderLayer, Transformer, VallinaTransformer, VanillaTransformer, apply_rotary_pos_emb,
    get_cos_sin, rotate_half)
