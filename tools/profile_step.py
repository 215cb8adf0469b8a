"""Kernel-time breakdown of one training step (torch.profiler, CUDA activities) for an L-layer slice of the headline
model.  Writes gpurun_out/profile_step_<tag>.txt.  Usage: python tools/profile_step.py [--layers 8] [--ops sm100]"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import bench  # noqa: E402

p = argparse.ArgumentParser()
p.add_argument("--layers", type=int, default=8)
p.add_argument("--ops", default="sm100")
p.add_argument("--tag", default="ours")
p.add_argument("--plain-gemm", default="native")
p.add_argument("--attention", default="auto")
p.add_argument("--preset", default="7b")
p.add_argument("--layout", default=None, help="tp,pp,dp")
p.add_argument("--sp", action="store_true")
a = p.parse_args()

from scaling_b200.core import LaunchConfig  # noqa: E402
from scaling_b200.ops import attention as ATT  # noqa: E402
from scaling_b200.ops import functional as OF  # noqa: E402
from scaling_b200.transformer.train import build_trainer  # noqa: E402

OF.set_ops(a.ops)
OF.set_plain_gemm(a.plain_gemm)
ATT.set_attention_impl(a.attention)
import os  # noqa: E402

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
args = argparse.Namespace(gpus=world, steps=1, warmup=1, layers=a.layers, ops=a.ops, act_ckpt="disabled", seqs_per_gpu=4, layout=a.layout,
                          preset=a.preset, sp=a.sp)
if "layout" in bench.PRESETS[a.preset]:
    bench.LAYOUTS[world] = bench.PRESETS[a.preset]["layout"](world)
if a.layout:
    bench.LAYOUTS[world] = tuple(int(v) for v in a.layout.split(","))
prefix = Path("/tmp/scaling_prof") / "data"
bench.write_synthetic_dataset(prefix, n_docs=64 * world, vocab=32000, rank=rank)
cfg = bench.model_config(args, world, prefix)
trainer = build_trainer(LaunchConfig(master_port=int(os.environ.get("MASTER_PORT", 29533)), master_addr="127.0.0.1",
                                     world_size=world, global_rank=rank, local_slot=rank, payload=cfg))
for _ in range(2):
    trainer.train_step()
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    out = trainer.train_step()
    torch.cuda.synchronize()
table = prof.key_averages().table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=90)
if rank == 0:
    Path("gpurun_out").mkdir(exist_ok=True)
    Path(f"gpurun_out/profile_step_{a.tag}.txt").write_text(table + f"\nstep_duration {out.step_duration}\n")
    print(table[-6000:])
    print("step_duration", out.step_duration)
if torch.distributed.is_initialized():
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()
