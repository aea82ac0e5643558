# 8-GPU validation + timing in one gpurun call
N=${1:-8}
export PYTHONPATH=.
mkdir -p gpurun_out
OUT=gpurun_out/multi_${N}gpu.txt
: > $OUT
say() { echo "$@" | tee -a "$OUT"; }
say "=== multirank + variant tests, $N GPUs"
timeout 1200 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_variants.py -q -m gpu -x 2>&1 | grep -v "^frame\|omitting" | tail -12 | tee -a $OUT
say "=== bench N=$N"
timeout 600 python bench.py --gpus $N --steps 5 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_${N}gpu.json | cut -c1-2500 | tee -a $OUT
if [ "$N" = "8" ]; then
say "=== bench N=4"
timeout 600 python bench.py --gpus 4 --steps 5 --warmup 3 --no-same-config 2>&1 | tail -1 | tee gpurun_out/bench_4gpu.json | cut -c1-2500 | tee -a $OUT
fi
for red in p2p nvls; do
  say "--- fit_probe reduce=$red N=$N"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29568 benchmarks/fit_probe.py --rows 100000000 --reps 3 --reduce $red 2>&1 | grep -v -i warn | tail -12 | tee -a $OUT
done
say "=== done"
