export PYTHONPATH=.
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | grep -v "^frame\|omitting" | tail -6
MTB_TIMING=1 timeout 600 python bench.py --gpus 1 --input f32 --steps 3 --warmup 3 --no-same-config --no-e2e 2>&1 | tail -3 | cut -c1-900
MTB_TIMING=1 timeout 600 python bench.py --gpus 2 --input f32 --steps 3 --warmup 3 --no-same-config --no-e2e 2>&1 | grep -v "^\[mtb timing\]" | tail -1 | cut -c1-900
