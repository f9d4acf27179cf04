set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_configs.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/t15_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/t15_pytest.log
timeout 300 python bench.py --no-cpu-baseline > gpurun_out/t15_bench.json 2>/dev/null; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/t15_bench.json')); print(d['ms_per_step'], {k:v['ms_avg'] for k,v in d['kernels'].items()})
for g,v in d['config']['grids'].items(): print(g, v['ms_per_step'])"
