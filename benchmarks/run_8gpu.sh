# 8-GPU validation + timing in one gpurun call
N=${1:-8}
export PYTHONPATH=.
mkdir -p gpurun_out
OUT=gpurun_out/multi_${N}gpu.txt
: > $OUT
say() { echo "$@" | tee -a "$OUT"; }
say "=== multirank tests, $N GPUs"
timeout 900 python -m pytest tests/test_gpu_multirank.py -q -m gpu -x -k "real_ranks or missing_rank" 2>&1 | grep -v "^frame\|omitting" | tail -6 | tee -a $OUT
say "=== bench N=$N"
timeout 600 python bench.py --gpus $N --steps 10 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_${N}gpu.json | cut -c1-2600 | tee -a $OUT
for M in 4 2; do
say "=== bench N=$M"
timeout 600 python bench.py --gpus $M --steps 10 --warmup 3 --no-same-config 2>&1 | tail -1 | tee gpurun_out/bench_${M}gpu.json | cut -c1-2300 | tee -a $OUT
done
say "=== bench f32 input N=$N"
timeout 600 python bench.py --gpus $N --input f32 --steps 5 --warmup 3 --no-same-config --no-e2e --no-nccl 2>&1 | tail -1 | tee gpurun_out/bench_${N}gpu_f32.json | cut -c1-900 | tee -a $OUT
say "=== configs 2 5 N=$N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29558 benchmarks/configs.py --cfg 2 5 2>&1 | grep "^{" | tee gpurun_out/configs_${N}gpu.jsonl | cut -c1-420 | tee -a $OUT
say "=== allreduce sweep N=$N (with the in-switch column)"
MTB_SWEEP_NVLS=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29566 benchmarks/allreduce_sweep.py 2>&1 | grep " B  nccl" | cut -c1-260 | tee gpurun_out/allreduce_sweep_${N}gpu.txt | tee -a $OUT
say "--- fit_probe reduce=p2p N=$N"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29568 benchmarks/fit_probe.py --rows 100000000 --reps 3 2>&1 | grep -v -i warn | tail -12 | tee -a $OUT
say "=== done"
