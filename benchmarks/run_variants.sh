# Development: time the 100M x 64 fit with each experimental library under mpitree_b200/lib/exp (tools/build_variant.sh).
export PYTHONPATH=.
for so in "" mpitree_b200/lib/exp/*.so; do
  echo "=== ${so:-default}"
  MTB_LIB=$so timeout 300 python benchmarks/fit_probe.py --rows 100000000 --reps 3 2>&1 | grep -E "rep 2|level  [139]" | cut -c1-250
done
