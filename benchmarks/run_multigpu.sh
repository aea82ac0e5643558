#!/bin/bash
# Multi-GPU validation + measurements in one gpurun call:  bash benchmarks/run_multigpu.sh <N> [quick]
N=${1:-8}
export PYTHONPATH=.
mkdir -p gpurun_out
echo "=== pytest (real ranks, world=$N)"; timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu -k real_ranks 2>&1 | tail -5
echo "=== bench p2p N=$N"; timeout 400 python bench.py --gpus $N --steps 3 --warmup 3 --reduce p2p 2>&1 | tail -1 | tee gpurun_out/bench_${N}gpu_p2p.json | cut -c1-300
echo "=== bench nccl N=$N"; timeout 400 python bench.py --gpus $N --steps 3 --warmup 3 --reduce nccl --no-e2e 2>&1 | tail -1 | tee gpurun_out/bench_${N}gpu_nccl.json | cut -c1-300
echo "=== fit_probe p2p N=$N"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 benchmarks/fit_probe.py --rows 100000000 --reps 3 2>&1 | grep -v -i warn | tail -13 | tee gpurun_out/fit_probe_${N}gpu.txt
echo "=== fit_probe nccl N=$N"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29557 benchmarks/fit_probe.py --rows 100000000 --reps 3 --reduce nccl 2>&1 | grep -v -i warn | grep "^rep" | tee gpurun_out/fit_probe_${N}gpu_nccl.txt
echo "=== configs N=$N"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29558 benchmarks/configs.py --cfg 2 4 5 2>&1 | grep "^{" | cut -c1-420
if [ "$2" != "quick" ]; then
  echo "=== allreduce sweep N=$N"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29556 benchmarks/allreduce_sweep.py 2>&1 | grep " B  nccl" | tee gpurun_out/allreduce_sweep_${N}gpu.txt
fi
if [ "$N" = "8" ]; then
  echo "=== bench p2p N=4"; timeout 400 python bench.py --gpus 4 --steps 3 --warmup 3 --reduce p2p 2>&1 | tail -1 | tee gpurun_out/bench_4gpu_p2p.json | cut -c1-300
fi
