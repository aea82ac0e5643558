export PYTHONPATH=.
timeout 600 python -m pytest tests/test_gpu_multirank.py -x -q -m gpu -k "missing_rank or real_ranks" 2>&1 | grep -v "^frame\|omitting" | tail -15
