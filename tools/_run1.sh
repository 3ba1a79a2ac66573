timeout 200 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python bench.py --no-cpu-baseline --no-e2e --steps 10 --warmup 3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('default', d['value'], d['ms_per_step'])"
for v in $VARIANTS; do WC_LIB=$PWD/cuda-mapreduce_b200/libwc_exp_$v.so timeout 100 python bench.py --no-cpu-baseline --no-e2e --steps 10 --warmup 3 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$v', d['value'], d['ms_per_step'])"; done
WC_LIB=$PWD/cuda-mapreduce_b200/libwc_exp_prof.so timeout 100 python tools/prof_probe.py 2>&1 | tail -1
