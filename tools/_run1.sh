timeout 250 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
WC_TIMING=1 timeout 600 python bench.py --workload config5 --steps 2 --warmup 3 --no-cpu-baseline 2>gpurun_out/c5_err.log | tail -1 > gpurun_out/bench_config5.json; grep "wc_finish" gpurun_out/c5_err.log | tail -1; python -c "
import json; d=json.load(open('gpurun_out/bench_config5.json')); print(d['value'], d['e2e'])"
