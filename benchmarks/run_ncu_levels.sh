# One ncu capture (after the plain run exits 0): root launch + levels 1..4 of the second fit of a 20M-row probe.
export PYTHONPATH=.
mkdir -p gpurun_out
CMD="python benchmarks/fit_probe.py --rows 20000000 --reps 2"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_level_fused -s 10 -c 5 -o /tmp/prof_levels -f $CMD > gpurun_out/ncu_levels.log 2>&1
tail -3 gpurun_out/plain.log
ls -la /tmp/prof_levels.ncu-rep
python tools/ncu_summary.py /tmp/prof_levels.ncu-rep | tee gpurun_out/ncu_summary_levels.md
ncu -i /tmp/prof_levels.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:k_level_fused2 > /tmp/prof_levels_source.csv 2>/dev/null
python benchmarks/ncu_lines.py /tmp/prof_levels_source.csv > gpurun_out/ncu_move_lines.txt 2>&1
head -45 gpurun_out/ncu_move_lines.txt | cut -c1-220
sz=$(stat -c %s /tmp/prof_levels.ncu-rep); if [ "$sz" -lt 50000000 ]; then cp /tmp/prof_levels.ncu-rep gpurun_out/; fi
