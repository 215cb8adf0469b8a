mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "lora" > gpurun_out/r2_pytest_lora.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_pytest_lora.log
timeout 300 python tools/profile_step.py --layers 8 --preset 7b_lora --tag r2_n1_lora > gpurun_out/prof_n1_lora.log 2>&1; echo rc=$?; tail -2 gpurun_out/prof_n1_lora.log
tools/round2_calls.sh n1
