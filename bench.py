#!/usr/bin/env python
"""Headline benchmark: tokens/s for the 7B GQA/SwiGLU/RoPE transformer (BASELINE.json config 3).

    python bench.py --gpus N --steps K --warmup W                  # this framework (scaling_b200)
    python bench.py --impl reference --gpus N --steps K --warmup W # unmodified reference from baseline/_ref

For N > 1 launch under ``python -m torch.distributed.run --nproc-per-node N`` (one rank per GPU).
Layouts follow BASELINE.md §4.1: N=1 → tp1·pp1·dp1, N=2 → tp2, N=4 → tp2·pp2, N=8 → tp2·pp2·dp2, ZeRO-1,
bf16, seq 4096, micro-batch 1.  Weak scaling: global batch = 8·N sequences, so every GPU sees the same
amount of work at every N.  Data is synthetic (random token ids written once to a memory-map in the
framework's own on-disk format and read back through its data loader) and weights are random-init.

Rank 0 prints ONE JSON line.  Both arms are timed with CUDA events on the device, bracketed by a barrier and
``torch.cuda.synchronize()`` on both sides, and the elapsed time is the max over ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent
SEQ = 4096
LAYOUTS = {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}  # n_gpus -> (tp, pp, dp)
SEQS_PER_GPU = 8  # weak scaling: global batch = 8 * N sequences of 4096 tokens


# BASELINE.json configs other than the headline: architecture / training overrides and the layout they are defined on
PRESETS = {
    "7b": {},
    # config 2: examples/transformer_example/config.yml geometry, pure data parallel at every N
    "example": {"arch": {"vocab_size": 128000, "hidden_size": 256, "num_layers": 4, "num_attention_heads": 2,
                         "attention_num_kv_heads": None, "attention_qkv_in_one": True, "mlp_factor": 2.5,
                         "sequence_length": 64, "masked_softmax": {"kernel": "torch", "softmax_in_fp32": True, "scale": 1.0},
                         "dropout_embedding": 0.1, "dropout_attention_probs": 0.1, "dropout_after_attention": 0.1,
                         "dropout_after_mlp": 0.1},
                "seq": 64, "micro_batch": 2, "layout": lambda n: (1, 1, n)},
    # config 4: Llama-13B geometry with 8 KV heads, TP = 8, activation checkpointing
    "13b_tp8": {"arch": {"hidden_size": 5120, "num_layers": 40, "num_attention_heads": 40, "attention_num_kv_heads": 8,
                         "mlp_factor": 2.7}, "layout": lambda n: (n, 1, 1), "act_ckpt": "every_layer"},
    # config 5: headline geometry + LoRA (rank 64 on q, k, v, dense) + adapters, frozen base, TP4·DP2 on 8 GPUs
    "7b_lora": {"arch": {"lora_config": {"name": "lora", "rank": 64, "alpha": 64},
                         "adapter_config": {"name": "adapter", "attention_downsampling_factor": 0.25,
                                            "mlp_downsampling_factor": 0.25}},
                "training": {"finetune": True, "finetunable_parameters": ["lora", "adapter"]},
                "layout": lambda n: (min(n, 4), 1, max(1, n // 4))},
}


def parse_args() -> argparse.Namespace:
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=4)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", choices=["ours", "reference"], default="ours")
    # development knobs; any use of them is recorded in the JSON line and makes the number non-headline
    p.add_argument("--layers", type=int, default=32)
    p.add_argument("--ops", choices=["sm100", "torch"], default="sm100")
    p.add_argument("--act-ckpt", choices=["disabled", "every_layer", "every_pipe_stage"], default="disabled")
    p.add_argument("--seqs-per-gpu", type=int, default=SEQS_PER_GPU)
    p.add_argument("--layout", default=None, help="development: tp,pp,dp override of the headline layout")
    p.add_argument("--sp", action="store_true", help="development: Megatron sequence parallelism over the TP group")
    p.add_argument("--preset", default="7b", choices=sorted(PRESETS),
                   help="7b = the headline (BASELINE config 3); the others are BASELINE configs 2, 4 and 5 "
                        "(non-headline, recorded in BASELINE.md §4.2)")
    return p.parse_args()


def dist_env() -> tuple[int, int, int, str, int]:
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", os.environ.get("LOCAL_SLOT", "0")))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    addr = os.environ.get("MASTER_ADDR", "127.0.0.1")
    port = int(os.environ.get("MASTER_PORT", "29513"))
    return rank, local, world, addr, port


def model_config(args: argparse.Namespace, n: int, data_prefix: Path) -> dict:
    """Config dict understood by both TransformerConfig classes (same field names by design)."""
    tp, pp, dp = LAYOUTS[n]
    gb = args.seqs_per_gpu * n
    cfg = _base_config(args, n, data_prefix, tp, pp, dp, gb)
    preset = PRESETS[getattr(args, "preset", "7b")]
    cfg["transformer_architecture"].update(preset.get("arch", {}))
    cfg["training"].update(preset.get("training", {}))
    if "act_ckpt" in preset and args.act_ckpt == "disabled":
        cfg["topology"]["activation_checkpointing_type"] = preset["act_ckpt"]
    if "micro_batch" in preset:
        mb = preset["micro_batch"]
        cfg["topology"]["micro_batch_size"] = mb
        cfg["topology"]["gradient_accumulation_steps"] = max(1, gb // (dp * mb))
    cfg["topology"]["sequence_parallel"] = bool(getattr(args, "sp", False))
    return cfg


def _base_config(args, n, data_prefix, tp, pp, dp, gb) -> dict:
    return {
        "logger": {"log_level": "warning", "log_dir": None, "use_wandb": False, "use_tensorboard": False},
        "topology": {
            "model_parallel_size": tp,
            "pipe_parallel_size": pp,
            "data_parallel_size": dp,
            "global_batch_size": gb,
            "micro_batch_size": 1,
            "gradient_accumulation_steps": gb // dp,
            "pipe_partition_method": "balanced",
            "activation_checkpointing_type": args.act_ckpt,
            "sequence_parallel": False,
        },
        "optimizer": {
            "beta1": 0.9,
            "beta2": 0.95,
            "eps": 1e-8,
            "gradient_clipping": 1.0,
            "zero": True,
            "loss_scaler": {"enable": False},
        },
        "learning_rate_scheduler": {
            "learning_rate": 1e-4,
            "learning_rate_minimum": 1e-5,
            "learning_rate_decay_style": "cosine",
            "learning_rate_decay_iters": 1000,
            "learning_rate_warmup_steps": 10,
        },
        "training": {"weight_decay": 0.1},
        "trainer": {
            "save_dir": None,
            "save_interval": None,
            "load_dir": None,
            "train_iterations": 10**6,
            "assert_checkpoint_loaded": False,
            "seed": 42,
            "dataloader_num_workers": 0,
            "dataloader_pin_memory": True,
            "delete_past_optimizer_states": False,
        },
        "profiler": {"profile_steps": 0},
        "transformer_architecture": {
            "vocab_size": 32000,
            "hidden_size": 4096,
            "num_layers": args.layers,
            "num_attention_heads": 32,
            "attention_num_kv_heads": 8,
            "attention_qkv_in_one": False,
            "rotary_embedding_base": 10000,
            "sequence_length": SEQ,
            "norm_type": "rms",
            "relative_position_embedding_type": "rotary_complex",
            "mlp_type": "swiglu",
            "mlp_factor": 3.5,
            "attention_bias": False,
            "mlp_bias": False,
            "weight_tying": False,
            "masked_softmax": {"kernel": "flash_attention", "softmax_in_fp32": True, "scale": 1.0},
            "layernorm": {"optimization_type": "torch", "layernorm_epsilon": 1e-5},
            "precision": "bfloat16",
            "dropout_embedding": 0.0,
            "dropout_attention_probs": 0.0,
            "dropout_after_attention": 0.0,
            "dropout_after_mlp": 0.0,
        },
        "data": {"data_prefixes": [str(data_prefix)], "use_mmap": True},
    }


def write_synthetic_dataset(prefix: Path, n_docs: int, vocab: int, rank: int) -> None:
    """Random-token memory map ({prefix}.bin/.idx/.meta.json, the format both frameworks read).

    Every document is exactly SEQ+1 tokens of ids in [1, vocab) (0 is the end-of-document id), so each training
    item is one full 4096-token causal sequence. Rank 0 writes, the others wait for the marker file.
    """
    import numpy as np

    ready = Path(str(prefix) + ".ready")
    if rank == 0:
        prefix.parent.mkdir(parents=True, exist_ok=True)
        for f in prefix.parent.glob(prefix.name + "*"):
            f.unlink()
        rng = np.random.default_rng(1234)
        doc_len = SEQ + 1
        with open(str(prefix) + ".bin", "wb") as fb, open(str(prefix) + ".idx", "wb") as fi:
            for i in range(n_docs):
                fb.write(rng.integers(1, vocab, size=doc_len, dtype=np.int32).tobytes())
                fi.write(np.array([i * doc_len, doc_len], dtype=np.int64).tobytes())
        json.dump(
            {"dtype": "int32", "index_dtype": "int64", "document_count": n_docs},
            open(str(prefix) + ".meta.json", "w"),
        )
        ready.write_text("ok")
    else:
        t0 = time.time()
        while not ready.is_file():
            time.sleep(0.2)
            if time.time() - t0 > 600:
                raise TimeoutError("synthetic dataset was not written by rank 0")


class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index: int):
        self.index = index
        self.samples: list[int] = []
        self.reasons: set[str] = set()
        self.max_mhz = 0
        self._stop = threading.Event()
        self._thread: threading.Thread | None = None

    def start(self) -> None:
        try:
            import pynvml

            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            return
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    def _run(self) -> None:
        nv = self._nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.2)

    def stop(self) -> dict:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=2)
        s = sorted(self.samples)
        return {
            "sm_mhz": s[len(s) // 2] if s else None,
            "sm_max_mhz": self.max_mhz or None,
            "reasons": sorted(self.reasons),
            "samples": len(s),
        }


def timed_steps(step_fn, warmup: int, steps: int, world: int, local: int):
    """W untimed steps, then exactly K steps between barrier+synchronize, CUDA events, max over ranks."""
    import torch
    import torch.distributed as dist

    def fence() -> None:
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    last = None
    for _ in range(warmup):
        last = step_fn()
    fence()
    sampler = ClockSampler(local)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        last = step_fn()
    e1.record()
    fence()
    clocks = sampler.stop()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), clocks, last


def mfu_block(args, tokens_per_s: float, n: int) -> dict:
    """PaLM model-FLOP utilisation, `tokens/s * (6 N + 12 L h s) / (peak * n_gpus)` with N = unique parameters
    (the reference's definition, src/scaling/transformer/utils/get_tflops.py:365-401, which itself divides by zero on
    B200).  Reported against the vendor dense bf16 peak and against the driver-measured cuBLAS sustained rate."""
    h, L, f, V, kvw = 4096, args.layers, 14336, 32000, 8 * 128
    n_params = L * (2 * h * h + 2 * h * kvw + 3 * h * f + 2 * h) + 2 * V * h + h
    flop_per_token = 6 * n_params + 12 * L * h * SEQ
    achieved = tokens_per_s * flop_per_token / n
    out = {"definition": "palm 6N+12Lhs, N=unique params", "params": n_params,
           "tflops_per_gpu": achieved / 1e12, "vs_vendor_peak_2250": achieved / 2.25e15}
    try:
        peaks = json.load(open(REPO / "MEASURED_PEAKS.json"))
        out["vs_measured_sustained"] = achieved / (float(peaks["bf16_tflops_sustained"]) * 1e12)
        out["measured_sustained_tflops"] = float(peaks["bf16_tflops_sustained"])
    except Exception:  # noqa: BLE001
        out["vs_measured_sustained"] = None
    return out


def result_line(args, n, impl, ms_total, clocks, extra: dict) -> dict:
    tp, pp, dp = LAYOUTS[n]
    gb = args.seqs_per_gpu * n
    tokens = gb * SEQ * args.steps
    value = tokens / (ms_total / 1e3)
    line = {
        "metric": "tokens/sec (whole job), 7B GQA/SwiGLU/RoPE transformer, ZeRO-1 bf16, seq 4096",
        "value": value,
        "unit": "tokens/s",
        "per_gpu": value / n,
        "n_gpus": n,
        "steps": args.steps,
        "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps,
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,  # the reference publishes no numbers
        "dtype": "bf16",
        "data": "synthetic random tokens (memory-map), random-init weights",
        "impl": impl,
        "clocks": clocks,
        "config": {
            "model": f"mistral-7b geometry h4096 L{args.layers} 32q/8kv heads ffn14336 V32000",
            "global_batch": gb,
            "seq_len": SEQ,
            "micro_batch": 1,
            "grad_accum": gb // dp,
            "parallelism": f"tp{tp}pp{pp}dp{dp}",
            "zero": 1,
            "activation_checkpointing": args.act_ckpt,
            "l2": "inputs larger than L2 (14.5 GB of bf16 weights stream every step; no explicit flush)",
        },
    }
    line["mfu"] = mfu_block(args, value, n)
    line.update(extra)
    if args.layers != 32 or args.seqs_per_gpu != SEQS_PER_GPU or args.layout or args.sp or args.preset != "7b":
        line["non_headline"] = "development knobs changed the model/batch; not the BASELINE headline config"
        line["preset"] = args.preset
        line["sequence_parallel"] = bool(args.sp)
        if args.preset != "7b":
            line.pop("mfu", None)  # the MFU block is written for the headline geometry
            line["config"]["model"] = f"preset {args.preset}"
    return line


# --------------------------------------------------------------------------------------------------------------
# reference arm: unmodified reference from baseline/_ref through its own public API and stock code path
# --------------------------------------------------------------------------------------------------------------
def run_reference(args: argparse.Namespace) -> None:
    rank, local, world, addr, port = dist_env()
    ref_root = REPO / "baseline" / "_ref"
    if not (ref_root / "scaling" / "__init__.py").is_file():
        # (re)create the install from a reachable reference checkout: rank 0 installs, the others wait for it
        sys.path.insert(0, str(REPO / "baseline"))
        from install_reference import ensure_reference

        why = ""
        if local == 0:
            ok, why = ensure_reference()
        t0 = time.time()
        while not (ref_root / "scaling" / "__init__.py").is_file() and local != 0 and time.time() - t0 < 120:
            time.sleep(0.5)
        if not (ref_root / "scaling" / "__init__.py").is_file():
            if rank == 0:
                print(json.dumps({"impl": "reference", "unavailable": why or "baseline/_ref is not installed"}))
            return
    sys.path.insert(0, str(ref_root))
    for name in [m for m in sys.modules if m == "scaling" or m.startswith("scaling.")]:
        del sys.modules[name]
    os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
    try:
        import torch
        import torch.utils._pytree as _pytree

        # Environment shim, outside the reference tree: the reference targets torch 2.1, where pytree TreeSpec objects
        # were mutable; its pipeline communicator assigns `spec.num_leaves` (communicator.py:186).  torch 2.11 froze
        # the dataclass, so give it back a plain __setattr__.  The reference source itself stays unmodified.
        if getattr(getattr(_pytree.TreeSpec, "__dataclass_params__", None), "frozen", False):
            for _cls in [_pytree.TreeSpec, *_pytree.TreeSpec.__subclasses__()]:  # LeafSpec has its own frozen setter
                if "__setattr__" in _cls.__dict__:
                    _cls.__setattr__ = lambda self, k, v: object.__setattr__(self, k, v)  # type: ignore
        from scaling.core import Topology  # type: ignore
        from scaling.core.logging import logger  # type: ignore
        from scaling.core.runner import LaunchConfig  # type: ignore
        from scaling.transformer import TransformerConfig, TransformerContext  # type: ignore
        from scaling.transformer.model import init_model, init_optimizer  # type: ignore
        from scaling.transformer.model.model import loss_function, metrics_aggregation_fn  # type: ignore
        from scaling.transformer.train import TransformerTrainer, _get_sync_batch, _read_datasets  # type: ignore
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": f"import failed: {type(e).__name__}: {e}"[:300]}))
        return

    n = args.gpus
    assert world == n, f"WORLD_SIZE {world} != --gpus {n}"
    gb = args.seqs_per_gpu * n
    prefix = Path("/tmp/scaling_bench_ref") / "data"
    write_synthetic_dataset(prefix, n_docs=gb * (args.steps + args.warmup + 2), vocab=32000, rank=rank)
    cfg = model_config(args, n, prefix)
    launch = LaunchConfig(
        master_port=port, master_addr=addr, world_size=world, global_rank=rank, local_slot=local, payload=cfg
    )
    try:
        config_dict = launch.overwrite_config_dict_with_launcher_args(cfg)
        config = TransformerConfig.from_dict(config_dict)
        topology = Topology(config=config.topology)
        logger.configure(config=config.logger, name=f"RANK {rank}", global_rank=rank)
        context = TransformerContext(config=config, topology=topology)
        context.initialize(master_addr=addr, master_port=str(port), seed=config.trainer.seed)
        model = init_model(context=context)
        optimizer = init_optimizer(context=context, model=model)
        dataset = None
        if topology.is_io_rank:
            dataset, _ = _read_datasets(context.config)
        trainer = TransformerTrainer(
            config=context.config.trainer,
            context=context,
            parallel_module=model,
            optimizer=optimizer,
            dataset=dataset,
            sync_batch_to_model_parallel=_get_sync_batch(context.config.data),
            loss_function=loss_function,
            metrics_aggregation_fn=metrics_aggregation_fn,
        )
        ms, clocks, out = timed_steps(trainer.train_step, args.warmup, args.steps, world, local)
    except Exception as e:  # noqa: BLE001
        import traceback

        traceback.print_exc()
        print(json.dumps({"impl": "reference", "rank": rank,
                          "unavailable": f"{type(e).__name__}: {str(e)[:240]}"}), flush=True)
        sys.stderr.flush()
        # peers may be blocked in a collective with this rank: leave without the final barrier so the launcher
        # tears the job down instead of hanging
        os._exit(0 if world == 1 else 1)
    if rank == 0:
        tokens_per_s = gb * SEQ * args.steps / (ms / 1e3)
        # the reference's step always loads its batch from the host loader and returns the loss to the host,
        # so its device-timed number already is its end-to-end number
        h2d = gb * (SEQ + 1) * 8
        extra = {
            "e2e": {"value": tokens_per_s, "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 8},
            "gpu_launches": 0,
            "last_loss": getattr(out, "loss", None),
            "peak_mem_gb": torch.cuda.max_memory_allocated() / 1e9,
        }
        print(json.dumps(result_line(args, n, "reference", ms, clocks, extra)))


# --------------------------------------------------------------------------------------------------------------
# this framework
# --------------------------------------------------------------------------------------------------------------
def run_ours(args: argparse.Namespace) -> None:
    from scaling_b200.transformer.bench_support import run_bench

    run_bench(args, sys.modules[__name__])


def main() -> None:
    args = parse_args()
    assert args.gpus in LAYOUTS, f"--gpus must be one of {sorted(LAYOUTS)}"
    preset = PRESETS[args.preset]
    if "layout" in preset:
        LAYOUTS[args.gpus] = preset["layout"](args.gpus)
    if "seq" in preset:
        global SEQ
        SEQ = preset["seq"]
    if args.layout:
        tp, pp, dp = (int(v) for v in args.layout.split(","))
        assert tp * pp * dp == args.gpus
        LAYOUTS[args.gpus] = (tp, pp, dp)
    if args.impl == "reference":
        run_reference(args)
    else:
        sys.path.insert(0, str(REPO))
        run_ours(args)
    try:
        import torch.distributed as dist

        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        pass


if __name__ == "__main__":
    main()
