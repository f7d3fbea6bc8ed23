"""Sine position table (reference models/position_encoding.py:7-35, 66-79).

``pe[p, 0, 2i] = sin(p*w_i)``, ``pe[p, 0, 2i+1] = cos(p*w_i)``, ``w_i = exp(-2i*ln(1e4)/d)``;
p indexes the *concatenated* target/location/context token sequence.  The buffer exists for
state_dict compatibility; the native token-assembly kernel reads the same table from HBM.
"""
import math

import torch
from torch import nn


class PositionalEncoding(nn.Module):
    def __init__(self, d_model: int, max_len: int = 5000):
        super().__init__()
        position = torch.arange(max_len).unsqueeze(1)
        div_term = torch.exp(torch.arange(0, d_model, 2) * (-math.log(10000.0) / d_model))
        pe = torch.zeros(max_len, 1, d_model)
        pe[:, 0, 0::2] = torch.sin(position * div_term)
        pe[:, 0, 1::2] = torch.cos(position * div_term)
        self.register_buffer("pe", pe)

    def forward(self, x):
        """x: [B, d, S] -> [B, d, S] (same layout the reference returns)."""
        b, _, s = x.shape
        return self.pe[:s, 0].t().unsqueeze(0).expand(b, -1, -1)


class PositionalEmbedding(nn.Module):
    """Learned position table of the concatenated sequence ('v3'/'learned'; reference
    models/position_encoding.py:38-63): ``dropout(LayerNorm(pos_embed[0:S]))``.  The module carries the parameters
    (state_dict-compatible); the native engine evaluates the table once per weight version for inference.  Training
    with this variant is not built (the trainer fails loudly on the missing sine buffer)."""

    def __init__(self, embedding_dim, dropout=0.1, max_position_embeddings=5000):
        super().__init__()
        self.pos_embed = nn.Embedding(num_embeddings=max_position_embeddings, embedding_dim=embedding_dim)
        self.LayerNorm = nn.LayerNorm(embedding_dim)
        self.dropout = nn.Dropout(dropout)

    def forward(self, x):
        raise RuntimeError("the position table is evaluated by the native engine; this module only holds parameters")


def build_position_encoding(config):
    if config.position_embedding in ("v2", "sine"):
        print("Using sine/cosine positional encodings")
        return PositionalEncoding(config.hidden_dim, max_len=1024)
    if config.position_embedding in ("v3", "learned"):
        print("Using learned positional encodings")
        return PositionalEmbedding(config.hidden_dim, max_position_embeddings=1024, dropout=0.1)
    raise ValueError(f"not supported {config.position_embedding}")
