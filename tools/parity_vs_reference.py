"""Loss-curve and checkpoint parity against the unmodified reference.

Three single-GPU phases, each its own process (the two frameworks cannot share a torch.distributed world):
  1. reference trains ONE step from random init and saves `global_step1` (weights, optimizer state, context);
  2. reference resumes from that checkpoint and trains N more steps;
  3. scaling_b200 resumes from THE SAME reference checkpoint (proves the on-disk format is shared, including the
     optimizer state and the data position) and trains the same N steps on the same data.
The driver mode runs all three and writes the two loss curves + relative differences as JSON.

    python tools/parity_vs_reference.py --steps 40 --layers 4 --out gpurun_out/parity.json
"""
import argparse
import json
import os
import subprocess
import sys
from pathlib import Path

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))
import bench  # noqa: E402

WORK = Path("/tmp/scaling_parity")
SEQ = bench.SEQ


def write_learnable_dataset(prefix: Path, n_docs: int, vocab: int) -> None:
    """Documents of SEQ+1 tokens following next = (3·t + 1) mod (vocab − 1) + 1 from a random start: learnable, so the
    curves actually move and a wrong gradient shows."""
    import numpy as np

    prefix.parent.mkdir(parents=True, exist_ok=True)
    rng = np.random.default_rng(7)
    doc_len = SEQ + 1
    with open(str(prefix) + ".bin", "wb") as fb, open(str(prefix) + ".idx", "wb") as fi:
        for i in range(n_docs):
            t = np.empty(doc_len, dtype=np.int64)
            t[0] = rng.integers(1, vocab)
            mult = 3
            for k in range(1, doc_len):
                t[k] = (mult * t[k - 1] + 1) % (vocab - 1) + 1
            fb.write(t.astype(np.int32).tobytes())
            fi.write(np.array([i * doc_len, doc_len], dtype=np.int64).tobytes())
    json.dump({"dtype": "int32", "index_dtype": "int64", "document_count": n_docs}, open(str(prefix) + ".meta.json", "w"))


def config(args, trainer_over: dict) -> dict:
    ns = argparse.Namespace(gpus=1, steps=args.steps, warmup=0, layers=args.layers, ops="sm100", act_ckpt="disabled",
                            seqs_per_gpu=args.seqs, layout=None)
    cfg = bench.model_config(ns, 1, WORK / "data" / "tokens")
    cfg["trainer"].update(trainer_over)
    cfg["learning_rate_scheduler"].update(learning_rate=1e-4, learning_rate_warmup_steps=10)
    return cfg


def run_arm(args) -> None:
    import torch

    port = 29640 + {"reference_init": 0, "reference": 1, "ours": 2, "ours_init": 3}[args.arm]
    ckpt = WORK / args.ckpt
    if args.arm in ("reference_init", "ours_init"):
        over = {"save_dir": str(ckpt), "save_interval": 1, "train_iterations": 1}
    else:
        over = {"load_dir": str(ckpt), "load_optimizer_states": True, "load_context": True, "assert_checkpoint_loaded": True,
                "train_iterations": 10**6}
    cfg = config(args, over)
    if args.arm.startswith("reference"):
        sys.path.insert(0, str(REPO / "baseline" / "_ref"))
        from scaling.core import Topology  # type: ignore
        from scaling.core.logging import logger  # type: ignore
        from scaling.core.runner import LaunchConfig  # type: ignore
        from scaling.transformer import TransformerConfig, TransformerContext  # type: ignore
        from scaling.transformer.model import init_model, init_optimizer  # type: ignore
        from scaling.transformer.model.model import loss_function, metrics_aggregation_fn  # type: ignore
        from scaling.transformer.train import TransformerTrainer, _get_sync_batch, _read_datasets  # type: ignore

        launch = LaunchConfig(master_port=port, master_addr="127.0.0.1", world_size=1, global_rank=0, local_slot=0, payload=cfg)
        c = TransformerConfig.from_dict(launch.overwrite_config_dict_with_launcher_args(cfg))
        topo = Topology(config=c.topology)
        logger.configure(config=c.logger, name="RANK 0", global_rank=0)
        ctx = TransformerContext(config=c, topology=topo)
        ctx.initialize(master_addr="127.0.0.1", master_port=str(port), seed=c.trainer.seed)
        model = init_model(context=ctx)
        opt = init_optimizer(context=ctx, model=model)
        data, _ = _read_datasets(ctx.config)
        trainer = TransformerTrainer(config=ctx.config.trainer, context=ctx, parallel_module=model, optimizer=opt, dataset=data,
                                     sync_batch_to_model_parallel=_get_sync_batch(ctx.config.data),
                                     loss_function=loss_function, metrics_aggregation_fn=metrics_aggregation_fn)
    else:
        from scaling_b200.core import LaunchConfig
        from scaling_b200.transformer.train import build_trainer

        launch = LaunchConfig(master_port=port, master_addr="127.0.0.1", world_size=1, global_rank=0, local_slot=0, payload=cfg)
        trainer = build_trainer(launch)
    if args.arm in ("reference_init", "ours_init"):
        # not run_training(): the reference's metric logging divides by a zero peak on GPUs it does not know (B200)
        trainer.train_step()
        trainer.save_checkpoint()
        return
    losses = []
    for _ in range(args.steps):
        out = trainer.train_step()
        losses.append(float(out.loss))
    torch.cuda.synchronize()
    json.dump(losses, open(WORK / f"losses_{args.arm}_{args.ckpt}.json", "w"))
    print(args.arm, "losses", [round(x, 4) for x in losses[:5]], "...", [round(x, 4) for x in losses[-3:]], flush=True)


def main() -> None:
    p = argparse.ArgumentParser()
    p.add_argument("--arm", default=None, choices=[None, "reference_init", "reference", "ours", "ours_init"])
    p.add_argument("--ckpt", default="ckpt", help="checkpoint sub-directory under the work dir")
    p.add_argument("--reverse", action="store_true",
                   help="the other direction: scaling_b200 writes the checkpoint, the unmodified reference resumes from it")
    p.add_argument("--steps", type=int, default=40)
    p.add_argument("--layers", type=int, default=4)
    p.add_argument("--seqs", type=int, default=4)
    p.add_argument("--out", default="gpurun_out/parity.json")
    args = p.parse_args()
    if args.arm is not None:
        run_arm(args)
        return
    import shutil

    shutil.rmtree(WORK, ignore_errors=True)
    write_learnable_dataset(WORK / "data" / "tokens", n_docs=args.seqs * (args.steps + 4), vocab=32000)
    ckpt = "ckpt_ours" if args.reverse else "ckpt"
    for arm in (("ours_init", "reference", "ours") if args.reverse else ("reference_init", "reference", "ours")):
        cmd = [sys.executable, __file__, "--arm", arm, "--steps", str(args.steps), "--layers", str(args.layers), "--seqs",
               str(args.seqs), "--ckpt", ckpt]
        env = dict(os.environ)
        if arm.startswith("reference"):
            # environment shim for the torch-2.1-era reference: its torch.load calls predate the weights_only default
            env["TORCH_FORCE_NO_WEIGHTS_ONLY_LOAD"] = "1"
        r = subprocess.run(cmd, cwd=str(REPO), env=env, timeout=900)
        if r.returncode != 0:
            raise SystemExit(f"phase {arm} failed with exit code {r.returncode}")
    ref = json.load(open(WORK / f"losses_reference_{ckpt}.json"))
    ours = json.load(open(WORK / f"losses_ours_{ckpt}.json"))
    rel = [abs(a - b) / abs(a) for a, b in zip(ref, ours)]
    result = {"config": f"mistral-7b geometry, {args.layers} layers, seq {SEQ}, {args.seqs} sequences/step, bf16, lr 1e-4 (10 warm-up steps, cosine)",
              "protocol": ("scaling_b200 trains 1 step and saves; the UNMODIFIED reference and scaling_b200 both resume from "
                           "that checkpoint (weights, optimizer state, data position) and train the same steps"
                           if args.reverse else
                           "reference trains 1 step and saves; both frameworks resume from that checkpoint (weights, "
                           "optimizer state, data position) and train the same steps"),
              "checkpoint_written_by": "scaling_b200" if args.reverse else "reference",
              "steps": args.steps, "reference": ref, "scaling_b200": ours, "relative_difference": rel,
              "max_relative_difference": max(rel), "first_step_relative_difference": rel[0],
              "mean_relative_difference_last_10": sum(rel[-10:]) / len(rel[-10:])}
    Path(args.out).parent.mkdir(exist_ok=True, parents=True)
    json.dump(result, open(args.out, "w"), indent=1)
    print(json.dumps({k: v for k, v in result.items() if k not in ("reference", "scaling_b200", "relative_difference")}))
    print("reference ", [round(x, 3) for x in ref])
    print("ours      ", [round(x, 3) for x in ours])


if __name__ == "__main__":
    main()
