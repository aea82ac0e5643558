export PYTHONPATH=.
timeout 300 python benchmarks/host_overhead.py 2>&1 | grep -v -i warn | head -75
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 benchmarks/host_overhead.py 2>&1 | grep -v -i warn | head -75
