#!/bin/bash
# One ncu capture of the fused level kernel with a chosen tile layout (one GPU; one capture per gpurun call).
#   bash benchmarks/ncu_staged.sh interleaved     # or: planar
# The command is first run without the profiler and must exit 0; numbers printed under ncu are not timings.
LAYOUT=${1:-interleaved}
export PYTHONPATH=. MTB_TILE=$LAYOUT
mkdir -p gpurun_out
CMD="python benchmarks/fit_probe.py --rows 20000000 --reps 1"
echo "=== plain run ($LAYOUT)"; timeout 300 $CMD 2>&1 | grep -v -i warn | tail -4 || exit 1
[ "${PIPESTATUS[0]}" = "0" ] || { echo "plain run failed: no capture"; exit 1; }
echo "=== ncu --set full, levels 3-4 of k_level_fused2 ($LAYOUT)"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_level_fused2 -s 3 -c 2 \
  -o gpurun_out/prof_fused2_$LAYOUT -f $CMD > gpurun_out/ncu_$LAYOUT.log 2>&1
ls -la gpurun_out/prof_fused2_$LAYOUT.ncu-rep 2>/dev/null || tail -5 gpurun_out/ncu_$LAYOUT.log
# source page by CUDA line (read here or on the CPU box with the same two commands)
ncu -i gpurun_out/prof_fused2_$LAYOUT.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/prof_fused2_${LAYOUT}_source.csv 2>/dev/null \
  && python benchmarks/ncu_lines.py gpurun_out/prof_fused2_${LAYOUT}_source.csv 25 | tail -30 || true
