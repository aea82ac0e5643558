export PYTHONPATH=.
mkdir -p gpurun_out
CMD="python benchmarks/fit_probe.py --rows 20000000 --reps 1"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_level_fused2 -s 8 -c 1 -o gpurun_out/prof_l9 -f $CMD > gpurun_out/ncu_l9.log 2>&1
tail -3 gpurun_out/plain.log
ncu -i gpurun_out/prof_l9.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/prof_l9_source.csv 2>/dev/null
python benchmarks/ncu_lines.py gpurun_out/prof_l9_source.csv 45 | cut -c1-260
