export PYTHONPATH=.
mkdir -p gpurun_out
CMD="python benchmarks/fit_probe.py --rows 20000000 --reps 2"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_level_fused -s 10 -c 10 -o gpurun_out/prof_levels -f $CMD > gpurun_out/ncu_levels.log 2>&1
tail -3 gpurun_out/plain.log
ls -la gpurun_out/prof_levels.ncu-rep
python tools/ncu_summary.py gpurun_out/prof_levels.ncu-rep
ncu -i gpurun_out/prof_levels.ncu-rep --page source --csv --print-source cuda,sass > gpurun_out/prof_levels_source.csv 2>/dev/null
