export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | grep -v "^frame\|omitting" | tail -12
timeout 300 python benchmarks/host_overhead.py 2>&1 | grep -v -i warn | head -22 | cut -c1-180
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 benchmarks/host_overhead.py 2>&1 | grep -v -i warn | grep -A22 "^world" | cut -c1-180
timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 --no-same-config 2>&1 | tail -1 | cut -c1-1800
timeout 600 python bench.py --gpus 2 --steps 5 --warmup 3 --no-same-config 2>&1 | tail -1 | cut -c1-1800
