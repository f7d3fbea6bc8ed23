"""Run under torchrun with >= 2 GPUs (tests/test_train_gpu.py::test_native_nccl_data_parallel launches it):
the native bucketed, overlapped NCCL all-reduce of the training step against one torch.distributed all-reduce of the
whole gradient arena after the backward pass, equal to fp32 rounding (the backward pass itself sums some gradients
with atomics, so two runs of it differ in the last bits)."""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from conftest import build_product_model
    from retr_b200 import native, synthetic
    import copy
    model, _, cfg = build_product_model("CaptionGlobalLoc")
    model = copy.deepcopy(model)
    model.load_state_dict(synthetic.calibrate_frozen_bn(model.state_dict()))
    model.precision = "bf16"
    model = model.to(dev).train()
    B = 4
    data = synthetic.make_batch(B, shard=rank, train=True)
    samples = tuple(s.to(dev) for s in synthetic.pack(data))
    caps, masks = data["caps"].to(dev), data["cap_masks"].to(dev)
    t, g, loc = model._split_samples(samples)
    inputs = model._train_inputs(t, g, loc, caps[:, :-1], masks[:, :-1])
    tr = model.trainer(B)
    buckets = tr.grad_buckets()
    assert len(buckets) == 6, buckets
    spans = sorted(buckets)
    assert spans[0][0] == 0 and spans[-1][1] == tr.gflat.numel()
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:])), spans          # the buckets tile the arena

    # (a) one torch all-reduce after the pass
    tr.loss_and_grads(*inputs, caps[:, 1:], dropout_p=0.1, seed=77 + rank)
    local_grads = tr.gflat.clone()
    dist.all_reduce(tr.gflat)
    tr.gflat.div_(world)
    want = tr.gflat.clone()
    # (b) native: bucketed, on a side stream, overlapped with the backward pass
    assert tr.enable_data_parallel()
    tr.loss_and_grads(*inputs, caps[:, 1:], dropout_p=0.1, seed=77 + rank)
    torch.cuda.synchronize()
    got = tr.gflat.clone()
    # (c) explicit exchange through retr_allreduce_grads on gradients produced without the hook
    native.check(tr.lib.retr_train_set_comm(tr.handle, None, 1))
    tr.loss_and_grads(*inputs, caps[:, 1:], dropout_p=0.1, seed=77 + rank)
    # (the backward pass sums LayerNorm / embedding gradients with atomics: equal to rounding, not bit for bit)
    assert (tr.gflat - local_grads).abs().max().item() <= 2e-5 * local_grads.abs().max().item()
    native.check(tr.lib.retr_allreduce_grads(tr.handle, tr._comm, native.stream_ptr(dev)))
    torch.cuda.synchronize()
    got2 = tr.gflat.clone()
    native.check(tr.lib.retr_train_set_comm(tr.handle, tr._comm, world))
    tol = 2e-5 * want.abs().max().item()
    assert (got - want).abs().max().item() <= tol, (got - want).abs().max().item()
    assert (got2 - want).abs().max().item() <= tol, (got2 - want).abs().max().item()
    assert (want - local_grads).abs().max().item() > 100 * tol          # the exchange did change the gradients
    if rank == 0:
        print(f"DP_NCCL_OK world={world} buckets={[(hi - lo) for lo, hi in buckets]}")
    tr.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
