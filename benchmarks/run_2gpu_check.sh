export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -q -m gpu 2>&1 | grep -v "^frame\|omitting" | tail -6 | tee gpurun_out/pytest_gpu_2gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py --gpus 2 --steps 5 --warmup 3 --no-same-config --no-e2e 2>&1 | tail -1 | cut -c1-700
