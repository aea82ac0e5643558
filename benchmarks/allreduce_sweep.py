"""u32 histogram all-reduce sweep 4 KB - 256 MB: peer one-shot / two-shot vs ncclAllReduce.

``MTB_SWEEP_NVLS=1`` adds the in-switch variant (multimem.ld_reduce / multimem.st over a multicast
mapping; staged until it has run on hardware) as a fourth column.

Run under torchrun, one rank per GPU.  Times are CUDA-event device times, max over ranks;
bus bandwidth uses the usual all-reduce convention 2 (P-1)/P x bytes / time.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpitree_b200.parallel.allreduce import PeerAllReduce  # noqa: E402
from mpitree_b200.parallel.comm import TorchComm, init_process_group_from_env  # noqa: E402


def timed(fn, iters, dev):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    assert init_process_group_from_env("nccl")
    comm = TorchComm()
    rank, world = comm.rank, comm.size
    dev = torch.device("cuda", torch.cuda.current_device())
    max_bytes = int(os.environ.get("MTB_SWEEP_MAX", 256 << 20))
    use_nvls = os.environ.get("MTB_SWEEP_NVLS") == "1"
    ar = PeerAllReduce(comm, max_bytes // 4, nvls=use_nvls)
    rows = []
    size = 4 << 10
    while size <= max_bytes:
        n = size // 4
        x = torch.randint(0, 1000, (n,), dtype=torch.int32, device=dev)
        ar.input_view(n).copy_(x)
        want = x.clone()
        dist.all_reduce(want)
        ok = []
        for algo in (0, 1, 2) if use_nvls else (0, 1):
            ar.launch(n, algo)
            torch.cuda.synchronize()
            ok.append(bool(torch.equal(ar.result_view(n), want)))
        iters = 200 if size <= (1 << 20) else (50 if size <= (32 << 20) else 10)
        y = x.clone()
        t_nccl = timed(lambda: dist.all_reduce(y), iters, dev)
        t_one = timed(lambda: ar.launch(n, 0), iters, dev)
        t_two = timed(lambda: ar.launch(n, 1), iters, dev)
        t_nvls = timed(lambda: ar.launch(n, 2), iters, dev) if use_nvls else float("nan")
        bus = lambda ms: 2 * (world - 1) / world * size / (ms * 1e-3) / 1e9  # noqa: E731
        t_auto = timed(lambda: ar.launch(n, -1), iters, dev)
        rows.append({"bytes": size, "auto_us": t_auto * 1e3, "nccl_us": t_nccl * 1e3, "oneshot_us": t_one * 1e3, "twoshot_us": t_two * 1e3,
                     "nccl_busbw_GBs": bus(t_nccl), "oneshot_busbw_GBs": bus(t_one), "twoshot_busbw_GBs": bus(t_two),
                     "nvls_us": t_nvls * 1e3, "nvls_busbw_GBs": bus(t_nvls), "correct": all(ok)})
        if rank == 0:
            r = rows[-1]
            print(f"{size:>10d} B  nccl {r['nccl_us']:9.1f} us  one-shot {r['oneshot_us']:9.1f} us  two-shot {r['twoshot_us']:9.1f} us"
                  f"  busbw nccl/1s/2s = {r['nccl_busbw_GBs']:7.1f} {r['oneshot_busbw_GBs']:7.1f} {r['twoshot_busbw_GBs']:7.1f} GB/s  ok={r['correct']}"
                  + (f"  nvls {r['nvls_us']:9.1f} us {r['nvls_busbw_GBs']:7.1f} GB/s" if use_nvls else "")
                  + f"  auto {r['auto_us']:9.1f} us ({'<=' if r['auto_us'] <= r['nccl_us'] else '>'} nccl)",
                  flush=True)
        size *= 4
    if rank == 0:
        os.makedirs("gpurun_out", exist_ok=True)
        json.dump({"world": world, "rows": rows}, open(f"gpurun_out/allreduce_sweep_{world}gpu.json", "w"), indent=1)
    dist.barrier()
    ar.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
