"""``from models.model import Transformer`` (reference models/model.py)."""
from distributed_pytorch_from_scratch_b200.models.model import (  # noqa: F401
    FFN, Attention, DecoderLayer, Transformer, VallinaTransformer, VanillaTransformer, apply_rotary_pos_emb,
    get_cos_sin, rotate_half)
