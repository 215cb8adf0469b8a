#!/bin/bash
# Measurement campaign for one GPU count: both arms of the headline benchmark (BASELINE.md §2 protocol), run back to back
# on the same box.  usage: tools/campaign.sh <n_gpus> <steps> <warmup> [extra bench args for OUR arm only]
# Writes gpurun_out/campaign_n<N>_{reference,ours}.json (+ .err).
N=$1; K=$2; W=$3; shift 3
mkdir -p gpurun_out
run() {  # impl tag extra...
  impl=$1; tag=$2; shift 2
  if [ "$N" = "1" ]; then
    python bench.py --impl $impl --gpus 1 --steps $K --warmup $W "$@" > gpurun_out/campaign_n${N}_${tag}.json 2> gpurun_out/campaign_n${N}_${tag}.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29700 + RANDOM % 200)) \
      bench.py --impl $impl --gpus $N --steps $K --warmup $W "$@" > gpurun_out/campaign_n${N}_${tag}.json 2> gpurun_out/campaign_n${N}_${tag}.err
  fi
  echo "$tag rc=$?"; tail -c 700 gpurun_out/campaign_n${N}_${tag}.json
}
run reference reference
run ours ours "$@"
