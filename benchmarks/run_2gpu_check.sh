export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | grep -v "^frame\|omitting" | tail -8
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29568 benchmarks/fit_probe.py --rows 400000 --reps 3 2>&1 | grep -v -i warn | tail -11
timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 --no-same-config 2>&1 | tail -1 | tee gpurun_out/bench_1gpu.json | cut -c1-2200
timeout 600 python bench.py --gpus 2 --steps 5 --warmup 3 --no-same-config 2>&1 | tail -1 | tee gpurun_out/bench_2gpu.json | cut -c1-2200
