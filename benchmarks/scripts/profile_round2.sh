set -x
python benchmarks/sweep.py --paths auto --out gpurun_out/r02_sweep_shapes.jsonl > gpurun_out/r02_sweep_shapes.log 2>&1
python benchmarks/sweep.py --surface --out gpurun_out/r02_surface.jsonl > gpurun_out/r02_surface.log 2>&1
python bench.py --steps 200 --warmup 20 > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r02_bench_1gpu_short.json 2>/dev/null
python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r02_bench_reference.json 2>/dev/null
python bench.py --steps 10 --warmup 20 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 10 --warmup 20 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
python bench.py --steps 10 --warmup 20 --no-cpu-baseline > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fused_sm_tile -s 25 -c 2 -o gpurun_out/r02_sm_tile python bench.py --steps 10 --warmup 20 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ncu -i gpurun_out/r02_sm_tile.ncu-rep --page raw --csv > gpurun_out/r02_sm_tile_raw.csv 2>/dev/null
ncu -i gpurun_out/r02_sm_tile.ncu-rep --page details > gpurun_out/r02_sm_tile_details.txt 2>/dev/null
tail -2 gpurun_out/ncu_full.log
cat gpurun_out/r02_sweep_shapes.log | cut -c1-200
