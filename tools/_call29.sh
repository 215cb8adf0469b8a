mkdir -p gpurun_out
REFTESTS_VERBOSE=1 timeout 200 python tools/run_reference_tests.py --suite gpu --reference baseline/_ref_tests --timeout 120 --tb short \
  --targets tests/core/test_optimizer tests/transformer/test_backwards_compatibility.py \
  --out gpurun_out/reference_gpu_tests_optimizer.json > gpurun_out/reference_gpu_tests_optimizer.log 2>&1; echo "ref rc=$?"; head -60 gpurun_out/reference_gpu_tests_optimizer.log | cut -c1-250
timeout 200 python -m pytest tests -m gpu -q -x -k "adamw or optimizer or zero or smoke" > gpurun_out/r2_pytest_opt.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_opt.log
