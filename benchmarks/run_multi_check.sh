# Multi-GPU validation + timing in one gpurun call:  bash benchmarks/run_multi_check.sh <N>
N=${1:-2}
export PYTHONPATH=.
mkdir -p gpurun_out
OUT=gpurun_out/multi_${N}gpu.txt
: > $OUT
say() { echo "$@" | tee -a "$OUT"; }
say "=== multirank + variant tests, $N GPUs"
timeout 1200 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_variants.py -q -m gpu -x 2>&1 | grep -v "^frame\|omitting" | tail -25 | tee -a $OUT
say "=== bench N=1 (no e2e)"
timeout 600 python bench.py --gpus 1 --steps 3 --warmup 3 --no-e2e --no-same-config 2>&1 | tail -1 | tee gpurun_out/bench_1gpu_on_${N}box.json | cut -c1-2000 | tee -a $OUT
say "=== bench N=$N"
timeout 600 python bench.py --gpus $N --steps 3 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_${N}gpu.json | cut -c1-2500 | tee -a $OUT
for red in p2p nvls; do
  say "--- fit_probe reduce=$red N=$N"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29568 benchmarks/fit_probe.py --rows 100000000 --reps 3 --reduce $red 2>&1 | grep -v -i warn | tail -13 | tee -a $OUT
done
say "=== done"
