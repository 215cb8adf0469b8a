mkdir -p gpurun_out
cp baseline/_ref_tests/tests/transformer/files/llama2-tokenizer.json /tmp/none 2>/dev/null
REFTESTS_VERBOSE=1 timeout 330 python tools/run_reference_tests.py --suite gpu --reference baseline/_ref_tests --timeout 120 \
  --targets tests/transformer/test_backwards_compatibility.py tests/core/test_optimizer tests/core/test_nn/test_flash_attention.py tests/core/test_nn/test_inference_module.py tests/core/test_nn/test_parallel_embedding.py \
  --out gpurun_out/reference_gpu_tests_core.json > gpurun_out/reference_gpu_tests_core.log 2>&1; echo "ref rc=$?"; head -40 gpurun_out/reference_gpu_tests_core.log | cut -c1-250
timeout 300 python -m pytest tests -m gpu -q -x > gpurun_out/r2_pytest_gpu_1gpu_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu_1gpu_final.log
