mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -rs > gpurun_out/r2_pytest_gpu_1gpu_c.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2_pytest_gpu_1gpu_c.log
python bench.py --steps 4 --warmup 3 > gpurun_out/bench_n1_after_fp16.json 2>/dev/null; echo "bench rc=$?"; tail -c 600 gpurun_out/bench_n1_after_fp16.json
