#!/usr/bin/env python
"""Secondary benchmark (BASELINE.json configs[3]): teacher-forced training step, batch 128/GPU, data
parallel with an NCCL all-reduce of the flat gradient buffer.

    python bench_train.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--dropout P]

One step = engine.train_step_native: train-mode forward (dropout), fused cross-entropy, backward,
gradient all-reduce (N > 1), clip_grad_norm_ + AdamW — reference engine.py:64-83.  Prints one JSON line
(samples/s over all ranks, ms/step, achieved TFLOP/s against the 3 x 54.65 GFLOP/sample of SURVEY §8d,
per-phase kernel counts).  The headline metric of the repo is bench.py; this line is for config 4.
"""
import argparse
import contextlib
import io
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
GFLOP_PER_SAMPLE_STEP = 3 * 54.65


def cpu_reference_train(cfg, batch=4, steps=2):
    """The reference's train step on the host cores (oracle port of engine.py:64-83: autograd + clip + torch AdamW)."""
    import time
    from oracle import retr_oracle as O
    from retr_b200 import synthetic
    from retr_b200.models import caption
    torch.manual_seed(cfg.seed)
    with contextlib.redirect_stdout(io.StringIO()):
        model, _ = caption.build_model(cfg)
    sd = synthetic.calibrate_frozen_bn(model.state_dict())
    names = [n for n, p in model.named_parameters() if p.requires_grad]
    sd = {k: v.clone().requires_grad_(k in names) for k, v in sd.items()}
    opt = torch.optim.AdamW([{"params": [sd[n] for n in names if "backbone" not in n]},
                             {"params": [sd[n] for n in names if "backbone" in n], "lr": cfg.lr_backbone}],
                            lr=cfg.lr, weight_decay=cfg.weight_decay)
    data = synthetic.make_batch(batch, shard=0, train=True)
    enc = (data["t_img"], data["t_mask"], data["g_img"], data["g_mask"], data["loc"])
    times = []
    for i in range(steps + 1):
        t0 = time.perf_counter()
        loss = O.train_loss(sd, cfg, *enc, data["caps"], data["cap_masks"], drop_p=cfg.dropout)
        opt.zero_grad()
        loss.backward()
        torch.nn.utils.clip_grad_norm_([sd[n] for n in names], cfg.clip_max_norm)
        opt.step()
        times.append(time.perf_counter() - t0)
    per_step = sum(times[1:]) / steps
    return batch / per_step, per_step, torch.get_num_threads()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--dropout", type=float, default=0.1)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    import torch.distributed as dist
    from retr_b200 import engine as E
    from retr_b200 import synthetic
    from retr_b200.configuration import Config
    from retr_b200.models import caption
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    cfg = Config(use_global_features=True, use_location_features=True)
    torch.manual_seed(cfg.seed)
    with contextlib.redirect_stdout(io.StringIO()):
        model, _ = caption.build_model(cfg)
    model.load_state_dict(synthetic.calibrate_frozen_bn(model.state_dict()))
    model.precision = args.precision
    model.dropout_p = args.dropout
    model = model.to(dev).train()
    torch.manual_seed(cfg.seed + rank)                    # main.py:19: seed offset by rank (dropout streams differ)
    B = args.batch
    data = synthetic.make_batch(B, shard=rank, train=True)
    samples = tuple(s.to(dev) for s in synthetic.pack(data))
    caps, masks = data["caps"].to(dev), data["cap_masks"].to(dev)

    def step():
        return E.train_step_native(model, samples, caps, masks, max_norm=cfg.clip_max_norm, lr=cfg.lr,
                                   lr_backbone=cfg.lr_backbone, weight_decay=cfg.weight_decay)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    losses = []
    for _ in range(args.warmup):
        losses.append(step()[0])
    tr = model.trainer(B)
    l0 = tr.kernel_launches()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        losses.append(step()[0])
    e1.record()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms)
    if rank == 0:
        total = B * world * args.steps
        value = total / (ms / 1000.0)
        peak = 1390.0
        p = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(p):
            peak = json.load(open(p)).get("bf16_tflops_sustained", peak)
        tf = value * GFLOP_PER_SAMPLE_STEP / 1000.0 / world
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            v, ps, cores = cpu_reference_train(cfg)
            cpu = {"value": v, "unit": "samples/s", "cores": cores, "kind": "port",
                   "sample": f"oracle port of engine.py:64-83 (autograd, clip, torch AdamW), batch 4, 2 timed steps at {ps:.2f} s/step"}
        print(json.dumps({
            "cpu_baseline": cpu,
            "metric": "training samples/sec (teacher-forced step: fwd + CE + bwd + clip + AdamW)", "value": value,
            "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "dtype": args.precision,
            "data": "synthetic",
            "config": {"workload": "CaptionGlobalLoc ResNet-101 d256 6+6 train step (BASELINE.json configs[3])",
                       "batch_per_gpu": B, "dropout": args.dropout, "parallelism": f"dp{world}"},
            "gpu_launches": int(tr.kernel_launches() - l0),
            "loss_first_last": [float(losses[0]), float(losses[-1])],
            "roofline": {"bound": "tensor", "achieved": tf, "peak": peak, "unit": "TFLOP/s per GPU", "frac": tf / peak,
                         "note": "algorithmic 3 x 54.65 GFLOP/sample (SURVEY §8d)"},
            "max_mem_gb": torch.cuda.max_memory_allocated(dev) / 2 ** 30,
        }))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
