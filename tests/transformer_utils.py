"""Helpers for transformer end-to-end tests: a tiny learnable dataset and the per-rank worker."""
from __future__ import annotations

from pathlib import Path
from typing import Any, Dict

import numpy as np


def write_dataset(prefix: Path, n_docs: int = 64, vocab: int = 128, seed: int = 0) -> None:
    """Documents of random length whose tokens follow next = (3·tok + 1) mod (vocab-1) + 1, ending with EOD (0)."""
    from scaling_b200.core import MemoryMapDatasetBuilder

    rng = np.random.default_rng(seed)
    prefix.parent.mkdir(parents=True, exist_ok=True)
    with MemoryMapDatasetBuilder(prefix) as b:
        for _ in range(n_docs):
            n = int(rng.integers(20, 90))
            t = np.empty(n, dtype=np.int64)
            t[0] = rng.integers(1, vocab)
            for i in range(1, n):
                t[i] = (3 * t[i - 1] + 1) % (vocab - 1) + 1
            t[-1] = 0
            b.add(t)


def tiny_config(data_prefix: Path, mp: int = 1, pp: int = 1, dp: int = 1, gas: int = 1, iters: int = 6,
                precision: str = "float32", **over: Any) -> Dict[str, Any]:
    arch = {
        "vocab_size": 128, "hidden_size": 64, "num_layers": 2, "num_attention_heads": 4, "sequence_length": 32,
        "norm_type": "rms", "relative_position_embedding_type": "rotary_complex", "mlp_type": "swiglu", "mlp_factor": 2.0,
        "attention_bias": False, "mlp_bias": False, "weight_tying": False, "attention_qkv_in_one": False,
        "attention_num_kv_heads": 2, "precision": precision,
        "masked_softmax": {"kernel": "torch", "softmax_in_fp32": True},
    }
    arch.update(over.pop("arch", {}))
    cfg = {
        "logger": {"log_level": "warning"},
        "topology": {"model_parallel_size": mp, "pipe_parallel_size": pp, "data_parallel_size": dp, "micro_batch_size": 2,
                     "gradient_accumulation_steps": gas, **over.pop("topology", {})},
        "optimizer": {"zero": True, "gradient_clipping": 1.0, **over.pop("optimizer", {})},
        "learning_rate_scheduler": {"learning_rate": 0.01, "learning_rate_decay_style": "constant",
                                    **over.pop("learning_rate_scheduler", {})},
        "training": {"weight_decay": 0.01, **over.pop("training", {})},
        "trainer": {"train_iterations": iters, "assert_checkpoint_loaded": False, "seed": 42,
                    "dataloader_pin_memory": False, "delete_past_optimizer_states": False, **over.pop("trainer", {})},
        "transformer_architecture": arch,
        "data": {"data_prefixes": [str(data_prefix)], **over.pop("data", {})},
    }
    assert not over, f"unused overrides {over}"
    return cfg


def run_transformer(rank: int, world_size: int, master_port: int, config_dict: Dict[str, Any]):
    from scaling_b200.core import LaunchConfig
    from scaling_b200.transformer.train import main

    launch = LaunchConfig(master_port=master_port, master_addr="127.0.0.1", world_size=world_size, global_rank=rank,
                          local_slot=rank, payload=config_dict)
    metrics = main(launch, return_metrics=True)
    return [m["training/loss"] for m in (metrics or [])]
