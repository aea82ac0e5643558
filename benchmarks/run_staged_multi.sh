export PYTHONPATH=.
mkdir -p gpurun_out
N=2
OUT=gpurun_out/staged_${N}gpu.txt
: > $OUT
say() { echo "$@" | tee -a "$OUT"; }
say "=== staged + multirank tests, $N GPUs"
timeout 900 python -m pytest tests/test_gpu_variants.py tests/test_gpu_multirank.py -q -m gpu -x 2>&1 | tail -25 | tee -a $OUT
say "=== allreduce sweep with NVLS column"
MTB_SWEEP_NVLS=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29566 benchmarks/allreduce_sweep.py 2>&1 | grep -v -i warn | tail -30 | tee -a $OUT
for red in p2p nvls nccl; do
  say "--- fit_probe reduce=$red"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29568 benchmarks/fit_probe.py --rows 100000000 --reps 3 --reduce $red 2>&1 | grep -v -i warn | tail -13 | tee -a $OUT
done
say "=== done"
