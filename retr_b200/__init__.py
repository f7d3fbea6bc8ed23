"""retr_b200 — Blackwell-native (sm_100a) hot path of RE⫶TR referring-expression generation.

Python surface mirrors the reference (models.caption.build_model, eval_utils.decode.greedy,
engine.train_one_epoch); the arithmetic lives in ``csrc/`` behind the C ABI declared in
``include/retr_b200.h``.
"""
from .configuration import Config, headline_config  # noqa: F401

__version__ = "0.1.0"
