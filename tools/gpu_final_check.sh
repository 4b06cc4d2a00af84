python -m pytest tests -m gpu -x -q > gpurun_out/r5g_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -2 gpurun_out/r5g_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python bench.py > gpurun_out/r5g_bench.json 2> gpurun_out/r5g_bench.err; echo "bench rc=$?"
python -c "
import json
d=[json.loads(l) for l in open('gpurun_out/r5g_bench.json') if l.startswith('{')][-1]
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['fp64_pipe_util'], d['config']['parity_all_ok'], d['gpu_launches'])"
