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


def build_position_encoding(config):
    if config.position_embedding in ("v2", "sine"):
        print("Using sine/cosine positional encodings")
        return PositionalEncoding(config.hidden_dim, max_len=1024)
    if config.position_embedding in ("v3", "learned"):
        raise NotImplementedError("learned position embedding: SURVEY 8(f) item 4, not built yet")
    raise ValueError(f"not supported {config.position_embedding}")
