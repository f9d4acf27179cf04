set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "odd_sizes or handles_per_device or device_small_solve" > gpurun_out/t13_pytest.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/t13_pytest.log
