export PYTHONPATH=.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | grep -v "^frame\|omitting" | tail -25
echo "=== fit_probe 100M"
timeout 300 python benchmarks/fit_probe.py --rows 100000000 --reps 3 2>&1 | grep -v -i warn | tail -14
echo "=== bench"
timeout 400 python bench.py --steps 3 --warmup 3 2>&1 | tail -1 | cut -c1-600
echo "=== root pass from row-major bytes (inside the timed build), depth 1"
for m in direct off; do
MTB_ROOT=$m timeout 300 python benchmarks/fit_probe.py --rows 100000000 --reps 3 --root-input --depth 1 2>&1 | grep -E "rep 2|level  0" | cut -c1-200
done
