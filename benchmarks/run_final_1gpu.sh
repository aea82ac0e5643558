export PYTHONPATH=.
mkdir -p gpurun_out
timeout 600 python benchmarks/configs.py --cfg 2 4 5 2>&1 | grep "^{" | tee gpurun_out/configs_1gpu.jsonl | cut -c1-330
timeout 600 python bench.py --steps 10 --warmup 3 2>&1 | tail -1 | tee gpurun_out/bench_1gpu.json | cut -c1-1500
timeout 600 python bench.py --input f32 --steps 5 --warmup 3 --no-same-config --no-e2e 2>&1 | tail -1 | tee gpurun_out/bench_1gpu_f32.json | cut -c1-500
timeout 300 python benchmarks/fit_probe.py --rows 100000000 --reps 3 2>&1 | grep -v -i warn | tail -13 > gpurun_out/fit_probe_1gpu.txt
