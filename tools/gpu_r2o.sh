#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py -x -q > gpurun_out/r2o_pytest_conv.log 2>&1; echo "conv_exit=$?"; tail -5 gpurun_out/r2o_pytest_conv.log
timeout 900 python bench.py --steps 2 --warmup 3 --skip-cpu > gpurun_out/r2o_bench.json 2> gpurun_out/r2o_bench.err; echo "bench_exit=$?"; tail -3 gpurun_out/r2o_bench.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2o_bench.json').read().strip().splitlines()[-1])
print('nbody', d['value'], d['ms_per_step'], d['config']['kernel'])
for k,v in d['conv'].items():
    if 'value' in v: print(k, round(v['value']), 'e2e', round(v['e2e_GBps']), 'bound', v.get('e2e_pcie_bound_GBps'), v['roofline']['frac'])
print(d['conv'].get('pcie_probe'))
print({k:v['ok'] for k,v in d['parity_check'].items()})
PY
