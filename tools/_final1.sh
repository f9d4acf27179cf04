set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/f1_pytest.log 2>&1; echo "pytest rc=$?" > gpurun_out/f1_rc.txt
tail -4 gpurun_out/f1_pytest.log
timeout 900 python bench.py > gpurun_out/f1_bench.json 2> gpurun_out/f1_bench.err; echo "bench rc=$?" >> gpurun_out/f1_rc.txt
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/f1_bench_ref.json 2> gpurun_out/f1_bench_ref.err; echo "ref rc=$?" >> gpurun_out/f1_rc.txt
bash tools/ncu_r2.sh > gpurun_out/f1_ncu.log 2>&1; echo "ncu rc=$?" >> gpurun_out/f1_rc.txt
cat gpurun_out/f1_rc.txt
head -c 600 gpurun_out/f1_bench.json; echo
head -c 600 gpurun_out/f1_bench_ref.json; echo
