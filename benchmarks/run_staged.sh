#!/bin/bash
# Validate and time the staged (opt-in) paths of DESIGN.md section 9 in one gpurun call.
#
#   1 GPU :  bash benchmarks/run_staged.sh 1      (interleaved tile, quad float encode)
#   N GPUs:  bash benchmarks/run_staged.sh N      (+ NVLS region / all-reduce)
#
# Every step is wrapped in its own timeout; the correctness tests run first and a failing test
# skips the timing of that path (a number from a wrong kernel is worthless).
N=${1:-1}
export PYTHONPATH=.
mkdir -p gpurun_out
OUT=gpurun_out/staged_${N}gpu.txt
: > "$OUT"
say() { echo "$@" | tee -a "$OUT"; }

say "=== staged tests (MTB_STAGED=1), $N GPU(s)"
timeout 900 python -m pytest tests/test_gpu_variants.py -q -m gpu -x 2>&1 | tail -15 | tee -a "$OUT"
TESTS_OK=${PIPESTATUS[0]}

if [ "$TESTS_OK" = "0" ]; then
  say "=== fit_probe 100M x 64, planar vs interleaved tile (1 GPU, device-resident)"
  for layout in planar interleaved; do
    say "--- MTB_TILE=$layout"
    MTB_TILE=$layout timeout 300 python benchmarks/fit_probe.py --rows 100000000 --reps 3 2>&1 | grep -v -i warn | tail -14 | tee -a "$OUT"
  done
  say "=== cfg2 (10M x 32 float32), scalar vs quad encode"
  for enc in scalar quads; do
    say "--- MTB_ENCODE=$enc"
    MTB_ENCODE=$enc MTB_TIMING=1 timeout 300 python benchmarks/configs.py --cfg 2 2>&1 | grep -v -i warn | tail -12 | cut -c1-400 | tee -a "$OUT"
  done
  say "=== bench.py with the interleaved tile (compare with the default run of the same box)"
  timeout 400 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 | cut -c1-400 | tee -a "$OUT"
  MTB_TILE=interleaved timeout 400 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 | cut -c1-400 | tee -a "$OUT"
fi

if [ "$N" -gt 1 ] && [ "$TESTS_OK" = "0" ]; then
  say "=== allreduce sweep with the NVLS column, world=$N"
  MTB_SWEEP_NVLS=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
    --master-port 29566 benchmarks/allreduce_sweep.py 2>&1 | grep " B  nccl" | tee -a "$OUT"
  say "=== fit_probe world=$N, peer push vs in-switch reduction"
  for red in p2p nvls; do
    say "--- reduce=$red"
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
      --master-port 29568 benchmarks/fit_probe.py --rows 100000000 --reps 3 --reduce $red 2>&1 | grep -v -i warn | tail -13 | tee -a "$OUT"
  done
  say "=== fit_probe p2p world=$N, planar vs interleaved"
  for layout in planar interleaved; do
    say "--- MTB_TILE=$layout"
    MTB_TILE=$layout timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
      --master-port 29567 benchmarks/fit_probe.py --rows 100000000 --reps 3 2>&1 | grep -v -i warn | grep "^rep" | tee -a "$OUT"
  done
fi
say "=== done"
