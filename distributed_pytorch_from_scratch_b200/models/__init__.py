"""Model zoo: tensor-parallel layers, the LLaMA-style Transformer, and its single-device oracle."""
