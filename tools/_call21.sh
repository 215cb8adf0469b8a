mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 300 $T --master-port 29961 tools/profile_step.py --layers 8 --layout 4,1,1 --tag r2_n4_tp4_full > gpurun_out/prof_tp4_full.log 2>&1; echo rc=$?; tail -3 gpurun_out/prof_tp4_full.log
timeout 300 $T --master-port 29962 tools/profile_step.py --layers 8 --preset 7b_lora --layout 4,1,1 --tag r2_n4_tp4_lora > gpurun_out/prof_tp4_lora.log 2>&1; echo rc=$?; tail -3 gpurun_out/prof_tp4_lora.log
