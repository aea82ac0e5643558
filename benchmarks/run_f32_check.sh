export PYTHONPATH=.
timeout 900 python -m pytest tests/test_gpu_variants.py tests/test_gpu_engine.py -x -q -m gpu 2>&1 | grep -v "^frame\|omitting" | tail -5
for enc in slots; do
echo "--- MTB_ENCODE=$enc"
MTB_ENCODE=$enc MTB_TIMING=1 timeout 600 python bench.py --gpus 1 --input f32 --steps 3 --warmup 3 --no-same-config --no-e2e 2>&1 | tail -2 | cut -c1-330
done
