export PYTHONPATH=.
timeout 300 python benchmarks/fit_probe.py --rows 400000 --reps 3 2>&1 | grep -v -i warn | tail -13
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29568 benchmarks/fit_probe.py --rows 400000 --reps 3 2>&1 | grep -v -i warn | tail -13
