"""Concat-transformer parameter tree (reference models/ConcatTransformer.py and
models/transformer_modules.py).

Attribute names reproduce the reference's state_dict keys:
``encoder.layers.{l}.self_attn.{sublayer,norm}``, ``.ff.{sublayer.0,sublayer.2,norm}``,
``decoder.layers.{l}.{tgt_self_attn,tgt_src_cross_attn,ff}``, ``embeddings.*``, ``encoder.norm``,
``decoder.norm``, ``positional_encoding.pe``.  ``nn.MultiheadAttention``/``nn.Linear`` are used
as parameter containers only; the math runs in the native library (csrc/attention.cu,
csrc/gemm_*.cu, csrc/layernorm.cu).
"""
import copy

from torch import nn

from .position_encoding import build_position_encoding


class _PreNormBranch(nn.Module):
    """norm -> sublayer -> dropout -> +x wrapper (SelfAttResidual / CrossAttResidual / FFResidual,
    reference models/transformer_modules.py:14-97).  Holds ``sublayer``, ``dropout``, ``norm``."""

    def __init__(self, sublayer, dimension, dropout):
        super().__init__()
        self.sublayer = sublayer
        self.dropout = nn.Dropout(dropout)
        self.norm = nn.LayerNorm(dimension)


def _attention_branch(d, heads, p):
    return _PreNormBranch(nn.MultiheadAttention(d, heads, dropout=p), d, p)


def _ffn_branch(d, ff, p):
    return _PreNormBranch(nn.Sequential(nn.Linear(d, ff), nn.ReLU(), nn.Linear(ff, d)), d, p)


class EncoderLayer(nn.Module):
    def __init__(self, d, heads, ff, p):
        super().__init__()
        self.self_attn = _attention_branch(d, heads, p)
        self.ff = _ffn_branch(d, ff, p)


class DecoderLayer(nn.Module):
    def __init__(self, d, heads, ff, p):
        super().__init__()
        self.tgt_self_attn = _attention_branch(d, heads, p)
        self.tgt_src_cross_attn = _attention_branch(d, heads, p)
        self.ff = _ffn_branch(d, ff, p)


class LayerStack(nn.Module):
    def __init__(self, proto, n, norm):
        super().__init__()
        self.layers = nn.ModuleList(copy.deepcopy(proto) for _ in range(n))
        self.num_layers = n
        self.norm = norm


class DecoderEmbeddings(nn.Module):
    """LN_eps(E_word[id] + E_pos[t]) (+dropout in train) — reference
    models/transformer_modules.py:100-129."""

    def __init__(self, config):
        super().__init__()
        self.word_embeddings = nn.Embedding(config.vocab_size, config.hidden_dim,
                                            padding_idx=config.pad_token_id)
        self.position_embeddings = nn.Embedding(config.max_position_embeddings, config.hidden_dim)
        self.LayerNorm = nn.LayerNorm(config.hidden_dim, eps=config.layer_norm_eps)
        self.dropout = nn.Dropout(config.dropout)


class ConcatTransformer(nn.Module):
    def __init__(self, config, d_model=512, nhead=8, num_encoder_layers=6, num_decoder_layers=6,
                 dim_feedforward=2048, dropout=0.1, activation="relu", normalize_before=False):
        super().__init__()
        enc_proto = EncoderLayer(d_model, nhead, dim_feedforward, dropout)
        enc_norm = nn.LayerNorm(d_model) if normalize_before else None
        self.encoder = LayerStack(enc_proto, num_encoder_layers, enc_norm)
        self.positional_encoding = build_position_encoding(config)
        self.embeddings = DecoderEmbeddings(config)
        dec_proto = DecoderLayer(d_model, nhead, dim_feedforward, dropout)
        self.decoder = LayerStack(dec_proto, num_decoder_layers, nn.LayerNorm(d_model))
        # xavier on every >=2-D parameter incl. both embedding tables, so the padding_idx row
        # is NOT zero (reference models/ConcatTransformer.py:40-43)
        for p in self.parameters():
            if p.dim() > 1:
                nn.init.xavier_uniform_(p)
        self.d_model = d_model
        self.nhead = nhead
        self.dim_feedforward = dim_feedforward
        self.p_dropout = dropout


def build_transformer(config):
    return ConcatTransformer(
        config, d_model=config.hidden_dim, dropout=config.dropout, nhead=config.nheads,
        dim_feedforward=config.dim_feedforward, num_encoder_layers=config.enc_layers,
        num_decoder_layers=config.dec_layers, normalize_before=config.pre_norm)
