"""Offline checkpoint repack: reference ``.pth`` -> one flat, mmap-able weight file.

The reference saves ``{'epoch', 'model_state_dict', 'optimizer_state_dict', ...}`` with ``torch.save``
(train_utils/checkpoints.py:4-21) and evaluation loads ``checkpoint['model_state_dict']`` into a freshly built model
(eval_model.py:26-28).  ``pack_checkpoint`` validates such a file against the module tree ``build_model(config)``
produces (same keys and shapes as SURVEY Appendix B; torchvision's ``num_batches_tracked`` entries are dropped, as
``FrozenBatchNorm2d._load_from_state_dict`` does, models/backbone.py:31-39) and writes

    magic "RETRPACK1\\n" | uint64 header length | JSON header | 64-byte aligned fp32 tensors, state_dict order

whose header carries the Config attributes ``build_model`` reads, so ``load_packed`` rebuilds the model and binds its
weights to a memory map of the file (no pickle, no second copy in host memory); ``model.to('cuda')`` then streams them
through ``retr_load_weight``.  CLI:  python -m retr_b200.checkpoint in.pth out.retr [--global] [--location] [k=v ...]
"""
import json
import sys

import numpy as np
import torch

from .configuration import Config

MAGIC = b"RETRPACK1\n"
CONFIG_KEYS = ("backbone", "dilation", "lr_backbone", "position_embedding", "hidden_dim", "pad_token_id",
               "max_position_embeddings", "layer_norm_eps", "dropout", "vocab_size", "enc_layers", "dec_layers",
               "dim_feedforward", "nheads", "pre_norm", "use_global_features", "use_location_features")


def _build(config):
    import contextlib
    import io
    from .models import caption
    with contextlib.redirect_stdout(io.StringIO()):
        model, _ = caption.build_model(config)
    return model


def pack_checkpoint(src, dst, config):
    """src: path of a reference checkpoint, the loaded dict, or a bare state_dict."""
    ckpt = torch.load(src, map_location="cpu", weights_only=False) if isinstance(src, (str, bytes)) or hasattr(src, "read") else src
    sd = ckpt.get("model_state_dict", ckpt)
    sd = {k: v for k, v in sd.items() if not k.endswith("num_batches_tracked")}
    want = _build(config).state_dict()
    missing, extra = sorted(set(want) - set(sd)), sorted(set(sd) - set(want))
    if missing or extra:
        raise KeyError(f"checkpoint does not match build_model(config): missing {missing[:5]}{'...' if len(missing) > 5 else ''}, "
                       f"unexpected {extra[:5]}{'...' if len(extra) > 5 else ''}")
    entries, off = [], 0
    for k, ref in want.items():
        t = sd[k]
        if tuple(t.shape) != tuple(ref.shape):
            raise ValueError(f"{k}: shape {tuple(t.shape)} in the checkpoint, {tuple(ref.shape)} in the model")
        off = (off + 63) // 64 * 64
        entries.append({"name": k, "shape": list(t.shape), "offset": off})
        off += t.numel() * 4
    header = json.dumps({"config": {k: getattr(config, k) for k in CONFIG_KEYS},
                         "epoch": ckpt.get("epoch") if isinstance(ckpt, dict) else None,
                         "entries": entries, "data_bytes": off}).encode()
    base = (len(MAGIC) + 8 + len(header) + 63) // 64 * 64
    with open(dst, "wb") as f:
        f.write(MAGIC + np.uint64(len(header)).tobytes() + header)
        f.write(b"\0" * (base - f.tell()))
        for e in entries:
            f.write(b"\0" * (base + e["offset"] - f.tell()))
            f.write(sd[e["name"]].detach().to(torch.float32).contiguous().numpy().tobytes())
    return len(entries), base + off


def read_header(path):
    with open(path, "rb") as f:
        if f.read(len(MAGIC)) != MAGIC:
            raise ValueError(f"{path} is not a retr_b200 packed checkpoint")
        n = int(np.frombuffer(f.read(8), np.uint64)[0])
        header = json.loads(f.read(n))
    return header, (len(MAGIC) + 8 + n + 63) // 64 * 64


def load_packed(path, **config_overrides):
    """-> (model, config): the caption model with its parameters bound to a read-only memory map of ``path``."""
    header, base = read_header(path)
    config = Config(**{**header["config"], **config_overrides})
    model = _build(config)
    blob = np.memmap(path, dtype=np.uint8, mode="r", offset=base, shape=(header["data_bytes"],))
    sd = {}
    for e in header["entries"]:
        n = int(np.prod(e["shape"])) if e["shape"] else 1
        arr = np.frombuffer(blob, dtype=np.float32, count=n, offset=e["offset"]).reshape(e["shape"])
        sd[e["name"]] = torch.from_numpy(arr)            # a view of the map; load_state_dict copies into the parameters
    model.load_state_dict(sd)
    return model.eval(), config


def main(argv):
    if len(argv) < 2:
        print(__doc__)
        return 2
    over = {}
    for a in argv[2:]:
        if a == "--global":
            over["use_global_features"] = True
        elif a == "--location":
            over["use_location_features"] = True
        elif "=" in a:
            k, v = a.split("=", 1)
            over[k] = json.loads(v) if v[:1] in "0123456789-tf[\"" else v
    n, size = pack_checkpoint(argv[0], argv[1], Config(**over))
    print(f"packed {n} tensors, {size} bytes -> {argv[1]}")
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1:]))
