export PYTHONPATH=.
timeout 600 python -m pytest tests/test_gpu_multirank.py tests/test_gpu_variants.py -x -q -m gpu 2>&1 | grep -v "^frame\|omitting" | tail -6
MTB_SWEEP_NVLS=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29566 benchmarks/allreduce_sweep.py 2>&1 | grep " B  nccl" | cut -c1-260
