"""Time one device-resident fit with a per-level breakdown (development probe, 1 GPU)."""
import argparse
import json
import time

import torch

from mpitree_b200.ops import cuda_engine
from mpitree_b200.ops.cpu_engine import SplitParams


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=10_000_000)
    ap.add_argument("--features", type=int, default=64)
    ap.add_argument("--classes", type=int, default=10)
    ap.add_argument("--depth", type=int, default=10)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--unfused", action="store_true")
    ap.add_argument("--informative", type=int, default=8)
    ap.add_argument("--reduce", default="auto")
    ap.add_argument("--root-input", "--root-cluster", dest="root_cluster", action="store_true",
                    help="hand the builder raw row-major bytes: the root pass (MTB_ROOT=direct|cluster) is timed inside the build")
    a = ap.parse_args()
    import os
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    comm = None
    if world > 1:
        from mpitree_b200.parallel.comm import TorchComm, init_process_group_from_env

        init_process_group_from_env("nccl")
        comm = TorchComm()
        total = a.rows
        a.rows = total * (rank + 1) // world - total * rank // world
    dev = torch.device("cuda", torch.cuda.current_device())
    g = torch.Generator(device=dev).manual_seed(rank)
    X = torch.randint(0, 256, (a.rows, a.features), dtype=torch.uint8, device=dev, generator=g)
    w = torch.randn(a.informative, device=dev, generator=g)
    score = (X[:, : a.informative].float() - 127.5) @ w + 40 * torch.randn(a.rows, device=dev, generator=g)
    sd = float((w.pow(2).sum() * (256 ** 2 - 1) / 12.0 + 1600.0).sqrt())
    qs = torch.distributions.Normal(0.0, sd).icdf(torch.linspace(0, 1, a.classes + 1, device=dev)[1:-1])
    y = torch.bucketize(score, qs).to(torch.uint8)
    del score
    params = SplitParams(max_depth=a.depth)
    data = cuda_engine.DeviceDataset.allocate(a.rows, a.features, a.classes, dev)
    out = []
    for rep in range(a.reps):
        data.cur = 0
        if not a.root_cluster:
            data.encode_u8(X)
            data.set_labels(y)
        torch.cuda.synchronize()
        b = cuda_engine.TreeBuilder(data, 256, params, fused=not a.unfused, profile=True, comm=comm,
                                    n_global=a.rows * world, reduce=a.reduce)
        if a.root_cluster:  # retile + labels + root histogram inside the timed build (one cluster kernel)
            b.root_input = (X, y)
        t0 = time.perf_counter()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        flat = b.build()
        e.record()
        torch.cuda.synchronize()
        ms = s.elapsed_time(e)
        out.append(ms)
        if rank == 0:
          print(f"rep {rep}: {ms:.2f} ms ({a.rows * world / ms / 1e3:.1f} M rows/s total) nodes={flat.n_nodes} wall={time.perf_counter() - t0:.3f}s")
    st = b.stats
    if rank != 0:
        return
    print(json.dumps({"FT": st.FT, "stages": st.stages, "smem": st.smem_bytes, "m_cap": st.m_cap,
                      "hist_MB": st.hist_bytes / 2**20, "launches": st.launches}))
    for lv, d in enumerate(st.level_ms):
        extra = " ".join(f"{k[:-3]}={v:.3f}" for k, v in d.items() if k.endswith("_ms") and k not in ("search_ms", "partition_ms"))
        print(f"  level {lv:2d} slots={d['slots']:5d} search={d['search_ms']:8.3f} ms partition={d['partition_ms']:8.3f} ms | {extra}")


if __name__ == "__main__":
    main()
