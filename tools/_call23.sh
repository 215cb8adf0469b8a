mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 300 $T --master-port 29962 tools/profile_step.py --layers 8 --preset 7b_lora --layout 4,1,1 --tag r2_n4_tp4_lora_b > gpurun_out/prof_tp4_lora_b.log 2>&1; echo rc=$?; tail -2 gpurun_out/prof_tp4_lora_b.log
timeout 400 python -m pytest tests/test_fused_tp_gpu.py -m gpu -q -x -k "lora or LoRA" > gpurun_out/r2_pytest_lora_tp.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest_lora_tp.log
tools/campaign.sh 4 10 5
SCALING_B200_P2P_GROUPS=1 timeout 300 $T --master-port 29952 bench.py --gpus 4 --steps 6 --warmup 3 > gpurun_out/campaign_n4_ours_p2pgroups.json 2>/dev/null; tail -c 400 gpurun_out/campaign_n4_ours_p2pgroups.json
