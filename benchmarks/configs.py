"""The BASELINE.json configurations beyond the headline (device-timed, max over ranks).

  cfg2  10M x 32 continuous (float32), 2 classes, max_depth=8
  cfg4  50M x 128 mixed categorical + continuous, max_depth=16 (exercises subtree hand-off)
  cfg5  batched predict throughput on a depth-12 tree (rows scaled by --predict-rows, 1B default
        split across ranks and processed in batches that fit memory)

Run with `python benchmarks/configs.py --cfg 2` on one GPU or under torchrun for N GPUs.
Each result is one JSON line on rank 0 and is appended to gpurun_out/configs_<N>gpu.jsonl.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpitree_b200.ops import cuda_engine  # noqa: E402
from mpitree_b200.parallel.comm import init_process_group_from_env  # noqa: E402
from mpitree_b200.tree import ParallelDecisionTreeClassifier  # noqa: E402


def timed_fit(clf, X, y, steps, warmup, world, dev):
    for _ in range(warmup):
        clf.fit(X, y)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = cuda_engine.KERNEL_LAUNCHES["count"]
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        clf.fit(X, y)
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t), cuda_engine.KERNEL_LAUNCHES["count"] - l0


def emit(rank, world, rec):
    if rank == 0:
        print(json.dumps(rec), flush=True)
        os.makedirs("gpurun_out", exist_ok=True)
        with open(f"gpurun_out/configs_{world}gpu.jsonl", "a") as fh:
            fh.write(json.dumps(rec) + "\n")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cfg", type=int, nargs="+", default=[2, 4, 5])
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--scale", type=float, default=1.0, help="scale row counts (development)")
    ap.add_argument("--predict-rows", type=int, default=1_000_000_000)
    a = ap.parse_args()
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        init_process_group_from_env("nccl")
    g = torch.Generator(device=dev).manual_seed(100 + rank)

    if 2 in a.cfg:
        rows_total = int(10_000_000 * a.scale)
        n = rows_total * (rank + 1) // world - rows_total * rank // world
        X = torch.randn(n, 32, device=dev, generator=g)
        y = ((X[:, 0] + 0.5 * X[:, 5] * X[:, 9] - 0.3 * X[:, 17] + 0.8 * torch.randn(n, device=dev, generator=g)) > 0).to(torch.int64)
        clf = ParallelDecisionTreeClassifier(max_depth=8, data_layout="sharded", device="cuda")
        ms, launches = timed_fit(clf, X, y, a.steps, a.warmup, world, dev)
        emit(rank, world, {"config": "10M x 32 continuous (f32, quantile-binned to 256), 2 classes, max_depth=8",
                           "n_gpus": world, "rows": rows_total, "ms_per_fit": ms, "rows_per_s": rows_total / ms * 1e3,
                           "nodes": clf.flat_tree_.n_nodes, "gpu_launches": launches, "steps": a.steps})
        del X, y

    if 4 in a.cfg:
        rows_total = int(50_000_000 * a.scale)
        n = rows_total * (rank + 1) // world - rows_total * rank // world
        # 64 continuous features (256 quantile codes) + 64 categorical features (4..64 levels), as bin codes
        X = torch.empty(n, 128, dtype=torch.uint8, device=dev)
        X[:, :64] = torch.randint(0, 256, (n, 64), dtype=torch.uint8, device=dev, generator=g)
        for j in range(64):
            X[:, 64 + j] = torch.randint(0, 4 + (j % 16) * 4, (n,), dtype=torch.uint8, device=dev, generator=g)
        score = (X[:, 3].float() - 128) / 64 + (X[:, 70] % 3 == 0).float() * 1.5 - (X[:, 90] > 10).float() + \
            ((X[:, 10] > 100) & (X[:, 100] < 5)).float() * 2 + 1.2 * torch.randn(n, device=dev, generator=g)
        y = torch.bucketize(score, torch.tensor([-1.0, 0.0, 1.0, 2.0], device=dev)).to(torch.uint8)
        for handoff in (None, 2048):  # None = automatic (level-wise while the histogram slots fit)
            clf = ParallelDecisionTreeClassifier(max_depth=16, min_samples_split=2, data_layout="sharded", device="cuda",
                                                 handoff_rows=handoff)
            ms, launches = timed_fit(clf, X, y, max(1, a.steps - 1), 1, world, dev)
            st = cuda_engine.fit_from_torch.last_stats
            emit(rank, world, {"config": "50M x 128 mixed categorical+continuous (bin codes), 5 classes, max_depth=16",
                               "handoff_rows": handoff, "n_gpus": world, "rows": rows_total, "ms_per_fit": ms,
                               "rows_per_s": rows_total / ms * 1e3, "nodes": clf.flat_tree_.n_nodes,
                               "handed_off_subtrees": st.handed_off, "hist_slots": st.m_cap, "gpu_launches": launches})
        del X, y

    if 5 in a.cfg:
        # depth-12 tree trained on 4M rows, then batched predict of `predict_rows` rows (64 u8 features)
        Xt = torch.randint(0, 256, (4_000_000, 64), dtype=torch.uint8, device=dev, generator=torch.Generator(device=dev).manual_seed(7))
        yt = ((Xt[:, 1].int() + Xt[:, 20].int() * 2 - Xt[:, 40].int()) // 100 + 3).clamp(0, 9).to(torch.uint8)
        clf = ParallelDecisionTreeClassifier(max_depth=12, data_layout="replicated" if world == 1 else "sharded", device="cuda")
        lo, hi = 4_000_000 * rank // world, 4_000_000 * (rank + 1) // world
        clf.fit(Xt[lo:hi], yt[lo:hi])
        total = int(a.predict_rows * a.scale)
        mine = total * (rank + 1) // world - total * rank // world
        batch = 200_000_000
        Xb = torch.randint(0, 256, (min(batch, mine), 64), dtype=torch.uint8, device=dev, generator=g)
        clf.predict(Xb[:1000])
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        done = 0
        while done < mine:
            k = min(Xb.shape[0], mine - done)
            out = clf.predict(Xb[:k])
            done += k
        e.record()
        torch.cuda.synchronize()
        t = torch.tensor([s.elapsed_time(e)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
        emit(rank, world, {"config": "batched predict, depth-12 tree, 64 u8 features (same 200M-row device batch re-read: > L2)",
                           "n_gpus": world, "rows": total, "ms": ms, "rows_per_s": total / ms * 1e3,
                           "tree_nodes": clf.flat_tree_.n_nodes, "tree_depth": clf.get_depth(),
                           "bytes_read_per_row": 64, "GBps": total * 68 / ms / 1e6})
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
