"""world_size-2 gloo test (CPU) of the sample-sharding + gather logic used for multi-GPU decode:
N-way sharded decode == single-process decode on pruned sequences."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT

V, EOS, BOS = 40, 2, 1


def make_samples(lo, hi):
    from retr_b200.models.utils import NestedTensor
    n = hi - lo
    img = torch.arange(lo, hi, dtype=torch.float32).view(n, 1, 1, 1).expand(n, 3, 4, 4).contiguous()
    return (NestedTensor(img, torch.zeros(n, 4, 4, dtype=torch.bool)),)


def stub_model(samples, caption, cap_mask):
    """Deterministic per-sample token script: sample s emits (s+t)%7+3 until step s%5+1, then EOS."""
    B, T = caption.shape
    sid = samples.tensors[:, 0, 0, 0].long()
    filled = int((~cap_mask[0]).sum())
    out = torch.zeros(B, T, V)
    for b in range(B):
        s = int(sid[b])
        tok = EOS if filled - 1 >= s % 5 + 1 else (s + filled) % 7 + 3
        out[b, filled - 1, tok] = 1.0
    return out


def _worker(rank, world, port, n, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from retr_b200.sharded_eval import sharded_greedy
    rows = sharded_greedy(make_samples, n, stub_model, batch_size=4, max_len=12, device="cpu", bos_token=BOS,
                          eos_token=EOS)
    q.put((rank, rows))
    dist.destroy_process_group()


def test_two_rank_sharded_decode_equals_single_process():
    from retr_b200.sharded_eval import shard_range, sharded_greedy
    n = 13                                   # ragged: 7 + 6
    assert shard_range(n, 0, 2) == (0, 7) and shard_range(n, 1, 2) == (7, 13)
    assert shard_range(3, 3, 4) == (3, 3)    # empty shard
    single = sharded_greedy(make_samples, n, stub_model, batch_size=5, max_len=12, device="cpu", bos_token=BOS,
                            eos_token=EOS)
    assert len(single) == n and all(len(r) == s % 5 + 1 for s, r in enumerate(single))
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29613, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0] == single and got[1] == single
