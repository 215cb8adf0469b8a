#!/bin/bash
# Bundled GPU-box commands of the round-2 measurement campaign (one gpurun call each).
#   tools/round2_calls.sh n1 | n2 | n4 | n8
set -u
mkdir -p gpurun_out
case "$1" in
n1)
  tools/campaign.sh 1 20 10
  python bench.py --preset example --gpus 1 --steps 20 --warmup 10 > gpurun_out/campaign_example_n1_ours.json 2>/dev/null; tail -c 400 gpurun_out/campaign_example_n1_ours.json
  python bench.py --impl reference --preset example --gpus 1 --steps 20 --warmup 10 > gpurun_out/campaign_example_n1_reference.json 2>/dev/null; tail -c 400 gpurun_out/campaign_example_n1_reference.json
  ;;
n2)
  timeout 1500 python -m pytest tests/test_fused_tp_gpu.py -m gpu -q -rs > gpurun_out/r2_pytest_gpu_2gpus_final.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r2_pytest_gpu_2gpus_final.log
  tools/campaign.sh 2 20 10
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29951 bench.py --gpus 2 --steps 8 --warmup 4 --sp > gpurun_out/campaign_n2_ours_sp.json 2>/dev/null; tail -c 500 gpurun_out/campaign_n2_ours_sp.json
  ;;
n4)
  tools/campaign.sh 4 10 5
  SCALING_B200_P2P_GROUPS=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29952 bench.py --gpus 4 --steps 8 --warmup 4 > gpurun_out/campaign_n4_ours_p2pgroups.json 2>/dev/null; tail -c 500 gpurun_out/campaign_n4_ours_p2pgroups.json
  ;;
n8)
  tools/campaign.sh 8 10 5
  for preset in 13b_tp8 7b_lora; do
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29953 bench.py --gpus 8 --steps 5 --warmup 3 --preset $preset > gpurun_out/campaign_${preset}_n8_ours.json 2> gpurun_out/campaign_${preset}_n8_ours.err; echo "$preset rc=$?"; tail -c 500 gpurun_out/campaign_${preset}_n8_ours.json
  done
  ;;
esac
