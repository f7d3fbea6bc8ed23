#!/usr/bin/env python
"""BASELINE.json configs[2]: an eval_model.py-shaped run — 10 000 synthetic samples greedy-decoded (max_len 128),
sharded by sample over the GPUs of one box (retr_b200.sharded_eval.sharded_greedy: contiguous blocks per rank, batches
of 256 streamed through eval_utils.decode.greedy_stream from HOST tensors, pruned id rows gathered at the end;
reference loop: engine.py:125-187).  Timed end to end on the host clock around the whole call (host staging, H2D,
decode, D2H, pruning, gather), barrier on both sides, max over ranks.

    python bench_eval.py [--samples 10000] [--batch 256] [--precision bf16]      (torchrun for N > 1)

Also checks that sharding does not change results: rank 0 re-decodes the first batch of the LAST rank's shard as a
batch of its own and compares the pruned ids with what that rank produced.  Host sample generation is not what is
measured: every rank serves zero-copy slices of 3 pre-generated pinned batches (sample i of the run = sample i % 768).
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=10000)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    args = ap.parse_args()
    rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("LOCAL_RANK", 0), ("WORLD_SIZE", 1)))
    import torch.distributed as dist
    import bench
    from retr_b200 import sharded_eval, synthetic
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    model, cfg = bench.build_model(args.precision)
    model.max_batch = args.batch
    model = model.to(dev)
    # 3 distinct batches + a copy of the first appended: samples [lo, hi) of the run are rows [lo % period, +hi-lo) of the
    # pool, always one contiguous (pinned, zero-copy) slice
    parts = [synthetic.make_batch(args.batch, shard=100 + k) for k in range(3)]
    pool = {k: torch.cat([p[k] for p in parts] + [parts[0][k]]).pin_memory() for k in parts[0]}
    period = 3 * args.batch

    def make_samples(lo, hi):
        o = lo % period
        return synthetic.pack({k: v[o:o + hi - lo] for k, v in pool.items()})

    def run(n):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rows = sharded_eval.sharded_greedy(make_samples, n, model, args.batch, 128, dev, 101, 102, gather=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        dt = torch.tensor([time.perf_counter() - t0], device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        return rows, float(dt)

    run(min(args.samples, 2 * args.batch * world))          # warm-up: contexts, graphs, NCCL
    rows, dt = run(args.samples)
    assert len(rows) == args.samples
    # sharding parity: the last rank's first batch, decoded alone on rank 0
    lo, hi = sharded_eval.shard_range(args.samples, world - 1, world)
    hi = min(hi, lo + args.batch)
    from retr_b200.eval_utils import decode
    with torch.no_grad():
        ids = decode.greedy([s.to(dev) for s in make_samples(lo, hi)], model, max_len=128, device=dev, bos_token=101,
                            eos_token=102)
    mine = decode.prune_cap_ids(ids.cpu().tolist(), clean=True, pad_token=0, bos_token=101, eos_token=102)
    same = mine == rows[lo:hi]
    if rank == 0:
        print(json.dumps({
            "metric": "referring expressions/sec, sharded evaluation run (greedy, max_len 128, host tensors in, pruned ids out)",
            "value": args.samples / dt, "unit": "expr/s", "n_gpus": world, "samples": args.samples, "seconds": dt,
            "batch": args.batch, "dtype": args.precision, "scaling": "strong",
            "config": {"workload": "BASELINE.json configs[2]: 10k synthetic samples, contiguous block per rank, "
                                   "greedy_stream per rank (H2D prefetch), all_gather_object of pruned rows",
                       "timed": "host clock around sharded_greedy incl. staging, H2D, decode, D2H, prune, gather"},
            "sharding_parity": {"checked_rows": hi - lo, "identical": bool(same),
                                "what": f"rank 0 decoded samples [{lo},{hi}) of rank {world - 1}'s shard as its own batch"},
        }))
    assert same, "sharded ids differ from the same samples decoded on another rank"
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
