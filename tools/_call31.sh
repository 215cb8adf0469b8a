mkdir -p gpurun_out
REFTESTS_VERBOSE=1 timeout 150 python tools/run_reference_tests.py --suite gpu --reference baseline/_ref_tests --timeout 100 --tb short \
  --targets tests/core/test_nn/test_local_attention.py tests/core/test_nn/test_flash_attention.py \
  --out gpurun_out/reference_gpu_tests_attention.json > gpurun_out/reference_gpu_tests_attention.log 2>&1; echo "ref rc=$?"; head -30 gpurun_out/reference_gpu_tests_attention.log | cut -c1-250
timeout 120 python bench.py --steps 2 --warmup 1 > gpurun_out/bench_n1_final_sanity.json 2>/dev/null; echo "bench rc=$?"; tail -c 300 gpurun_out/bench_n1_final_sanity.json
