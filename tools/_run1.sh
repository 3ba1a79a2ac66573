timeout 250 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/b11.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_v11_launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu11.log 2>&1
tail -1 gpurun_out/b11.log | cut -c1-300
wc -l gpurun_out/r1_v11_launches.csv
