"""Parameter-efficient finetuning end to end on CPU: bitfit, soft prompt, adapters, LoRA, finetunable token ids —
frozen parameters stay bit-identical to the base checkpoint, tuned ones move, tuned ones land in separate checkpoint
files, and non-strict loading obeys its regex lists (reference: tests/transformer/test_finetuning.py,
test_load_checkpoint_non_strict.py)."""
from pathlib import Path

import pytest
import torch

from tests.transformer_utils import run_transformer, tiny_config, write_dataset
from tests.utils import dist_launcher


@pytest.fixture(scope="module")
def base(tmp_path_factory):
    root = tmp_path_factory.mktemp("ft")
    prefix = root / "data" / "tokens"
    write_dataset(prefix)
    ckpt = root / "base"
    arch = {"attention_bias": True, "mlp_bias": True, "norm_type": "layernorm"}
    dist_launcher(run_transformer, 1, config_dict=tiny_config(prefix, iters=2, arch=arch,
                                                              trainer={"save_dir": str(ckpt), "save_interval": 2}))
    return prefix, ckpt, arch


def _layer_files(step_dir: Path):
    return sorted(p.name for p in step_dir.glob("model_state_layer_*.pt"))


def _load_all(step_dir: Path):
    out = {}
    for f in step_dir.glob("model_state_layer_*.pt"):
        for k, v in torch.load(f, map_location="cpu", weights_only=False).items():
            out[f"{f.name.split('_')[3]}:{k}"] = v
    return out


CASES = {
    "bitfit": dict(arch={"bitfit_bias_config": {"name": "tune"}}, pattern="bias_tune", sep="bias_tune",
                   unexpected=[r"\.bias$"]),  # the base run's plain biases have no counterpart in the bitfit model
    "softprompt": dict(arch={"softprompt_config": {"name": "sp", "n_tokens": 4}}, pattern="softprompt_sp", sep="softprompt_sp"),
    "adapter": dict(arch={"adapter_config": {"name": "ad", "attention_downsampling_factor": 0.25, "mlp_downsampling_factor": 0.25}},
                    pattern="adapter_ad", sep="adapter_ad"),
    "lora": dict(arch={"lora_config": {"name": "lo", "rank": 4, "alpha": 8}}, pattern="lora_modules", sep=None),
}


@pytest.mark.parametrize("kind", list(CASES))
def test_finetuning_freezes_base_and_trains_new_parameters(tmp_path, base, kind):
    prefix, ckpt, arch = base
    case = CASES[kind]
    a = dict(arch, **case["arch"])
    out_dir = tmp_path / "tuned"
    allowed = [case["pattern"]]
    cfg = tiny_config(prefix, iters=4, arch=a,
                      training={"finetune": True, "finetunable_parameters": [case["pattern"]]},
                      trainer={"load_dir": str(ckpt), "load_optimizer_states": False, "load_context": False,
                               "allowed_missing_keys_in_checkpoint": allowed,
                               "allowed_unexpected_keys_in_checkpoint": case.get("unexpected", []),
                               "save_dir": str(out_dir), "save_interval": 2},
                      learning_rate_scheduler={"learning_rate": 0.05, "learning_rate_decay_style": "constant"})
    losses = dist_launcher(run_transformer, 1, config_dict=cfg)[0]
    assert len(losses) == 4 and all(x == x for x in losses)
    base_w = _load_all(ckpt / "global_step2")
    mid, end = _load_all(out_dir / "global_step2"), _load_all(out_dir / "global_step4")
    tuned = [k for k in end if case["pattern"] in k]
    frozen = [k for k in end if case["pattern"] not in k]
    assert tuned and frozen
    for k in frozen:  # the pretrained weights must not move by a single bit
        if k in base_w:
            assert torch.equal(end[k], base_w[k]), k
    assert any(not torch.equal(mid[k], end[k]) for k in tuned), "tuned parameters did not change between steps 2 and 4"
    if case["sep"] is not None:  # tuned parameters are written to their own files next to the base layer files
        names = _layer_files(out_dir / "global_step4")
        assert any(n.endswith(f"_{case['sep']}.pt") for n in names), names
        sep_file = next(n for n in names if n.endswith(f"_{case['sep']}.pt"))
        keys = torch.load(out_dir / "global_step4" / sep_file, map_location="cpu", weights_only=False).keys()
        assert all(case["pattern"] in k for k in keys)


def test_finetunable_token_ids_only_move_their_rows(tmp_path, base):
    prefix, ckpt, arch = base
    out_dir = tmp_path / "tok"
    cfg = tiny_config(prefix, iters=3, arch=dict(arch, finetunable_token_ids=[3, 5]),
                      training={"finetune": True, "finetunable_parameters": ["embedding.weight"], "weight_decay": 0.0},
                      trainer={"load_dir": str(ckpt), "load_optimizer_states": False, "load_context": False,
                               "save_dir": str(out_dir), "save_interval": 3})
    dist_launcher(run_transformer, 1, config_dict=cfg)
    before = _load_all(ckpt / "global_step2")["0:embedding.weight"]
    after = _load_all(out_dir / "global_step3")["0:embedding.weight"]
    moved = (before != after).any(dim=1).nonzero().flatten().tolist()
    assert set(moved) <= {3, 5} and moved, moved


def test_non_strict_load(tmp_path, base):
    prefix, ckpt, arch = base
    a = dict(arch, adapter_config={"name": "ad", "attention_downsampling_factor": 0.25, "mlp_downsampling_factor": 0.25})

    def run(**trainer):
        cfg = tiny_config(prefix, iters=3, arch=a, trainer={"load_dir": str(ckpt), "load_optimizer_states": False,
                                                            "load_context": False, **trainer})
        return dist_launcher(run_transformer, 1, config_dict=cfg)[0]

    with pytest.raises(AssertionError, match="Missing keys in checkpoint"):
        run()  # the adapter weights are not in the base checkpoint
    assert len(run(allowed_missing_keys_in_checkpoint=["adapter_ad"])) == 3
    # the other direction: a checkpoint with extra keys loads only when they are allowed / ignored
    with_adapter = tmp_path / "with_adapter"
    cfg = tiny_config(prefix, iters=1, arch=a, trainer={"save_dir": str(with_adapter), "save_interval": 1})
    dist_launcher(run_transformer, 1, config_dict=cfg)

    def run_plain(**trainer):
        cfg = tiny_config(prefix, iters=2, arch=arch, trainer={"load_dir": str(with_adapter), "load_optimizer_states": False,
                                                               "load_context": False, **trainer})
        return dist_launcher(run_transformer, 1, config_dict=cfg)[0]

    with pytest.raises(AssertionError, match="Unexpected keys in checkpoint"):
        run_plain()
    assert len(run_plain(allowed_unexpected_keys_in_checkpoint=["adapter_ad"])) == 2
    assert len(run_plain(ignore_keys_in_checkpoint=["adapter_ad"])) == 2
