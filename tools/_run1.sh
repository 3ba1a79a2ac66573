set -x
python bench.py --no-cpu-baseline --steps 5 --warmup 3 > gpurun_out/b10.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:count_kernel --launch-skip 6 --launch-count 1 -o gpurun_out/r1_v11 -f python bench.py --no-cpu-baseline --no-e2e --steps 1 --warmup 3 > gpurun_out/ncu10.log 2>&1
tail -1 gpurun_out/b10.log
