"""Data-parallel gradient exchange of the training step on CPU: world_size-2 gloo.

The arithmetic of a step is native (GPU only); what is tested here is the host logic around it —
`_NativeTrainForward.backward` all-reduces the trainer's flat gradient buffer and hands autograd the MEAN over
ranks, and `engine.train_step_native` does the same before the fused clip + AdamW.  A stub trainer stands in
for the native one: its "gradient" of every parameter is (rank + 1) * ones, so the expected result is 1.5.
"""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class StubTrainer:
    def __init__(self, model, rank):
        self.trainable = [n for n, p in model.named_parameters() if p.requires_grad]
        sizes = [p.numel() for n, p in model.named_parameters() if p.requires_grad]
        self.gflat = torch.zeros(sum(sizes))
        self.grads, off = {}, 0
        for (n, p), k in zip([(n, p) for n, p in model.named_parameters() if p.requires_grad], sizes):
            self.grads[n] = self.gflat[off:off + k].view(p.shape)
            off += k
        self.rank, self.T, self.V = rank, 128, 16
        self.clipped = None

    def forward(self, *inputs, dropout_p=0.0, seed=0):
        return torch.zeros(inputs[0].shape[0], self.T, self.V)

    def backward(self, dlogits):
        self.gflat.fill_(float(self.rank + 1))
        return self.grads

    def loss_and_grads(self, *inputs, dropout_p=0.0, seed=0):
        self.gflat.fill_(float(self.rank + 1))
        return torch.tensor(float(self.rank))

    def clip_adamw(self, max_norm, lr, lr_backbone, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-4):
        self.clipped = self.gflat.clone()
        return self.gflat.norm()


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from conftest import build_product_model
    from retr_b200 import engine as E
    from retr_b200.models.caption import _NativeTrainForward
    model, _, _ = build_product_model("Caption")
    stub = StubTrainer(model, rank)
    model.trainer = lambda batch=1: stub
    model.train()
    inputs = (torch.zeros(2, 3, 8, 8), torch.zeros(2, 8, 8, dtype=torch.bool), None, None, None,
              torch.zeros(2, 128, dtype=torch.long), torch.zeros(2, 128, dtype=torch.bool))
    named = [(n, p) for n, p in model.named_parameters() if p.requires_grad]
    params = [p for _, p in named]
    out = _NativeTrainForward.apply(model, inputs, 0.0, 0, [n for n, _ in named], *params)
    out.sum().backward()
    ok_autograd = all(torch.allclose(p.grad, torch.full_like(p, 1.5)) for p in params)
    frozen_untouched = all(p.grad is None for p in model.parameters() if not p.requires_grad)
    # fused path
    model._split_samples = lambda samples: (samples[0], None, None)
    model._train_inputs = lambda *a: inputs
    E.train_step_native(model, (inputs[0],), torch.zeros(2, 129, dtype=torch.long), torch.zeros(2, 129, dtype=torch.bool),
                        max_norm=0.1, lr=1e-4, lr_backbone=1e-5)
    ok_fused = bool(torch.allclose(stub.clipped, torch.full_like(stub.clipped, 1.5)))
    q.put((rank, ok_autograd, frozen_untouched, ok_fused))
    dist.destroy_process_group()


def test_two_rank_gradient_mean():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29631, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert sorted(g[0] for g in got) == [0, 1]
    assert all(g[1] and g[2] and g[3] for g in got), got
