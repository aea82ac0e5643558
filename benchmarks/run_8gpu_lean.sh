# Lean N-GPU timing (no tests): headline bench with the NCCL-reduction baseline and the end-to-end arm, then the per-level breakdown.
N=${1:-8}
export PYTHONPATH=.
mkdir -p gpurun_out
OUT=gpurun_out/multi_${N}gpu_lean.txt
: > $OUT
echo "=== bench N=$N" | tee -a $OUT
timeout 600 python bench.py --gpus $N --steps 10 --warmup 3 --no-same-config 2>&1 | tail -1 | tee gpurun_out/bench_${N}gpu.json | cut -c1-2300 | tee -a $OUT
echo "--- fit_probe reduce=p2p N=$N" | tee -a $OUT
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29568 benchmarks/fit_probe.py --rows 100000000 --reps 3 2>&1 | grep -v -i warn | tail -12 | tee -a $OUT
