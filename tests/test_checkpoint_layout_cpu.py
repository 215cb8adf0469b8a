"""On-disk checkpoint contract: directory layout, file names, state-dict keys, optimizer-state structure
and layout independence (a TP=2 run and a single-process run write the same files with the same shapes)."""
from pathlib import Path

import pytest
import torch

from tests.transformer_utils import run_transformer, tiny_config, write_dataset
from tests.utils import dist_launcher


@pytest.fixture(scope="module")
def runs(tmp_path_factory):
    root = tmp_path_factory.mktemp("ckpt_layout")
    prefix = root / "data" / "tokens"
    write_dataset(prefix)
    out = {}
    for name, (mp, pp) in {"single": (1, 1), "tp2": (2, 1), "pp2": (1, 2)}.items():
        d = root / name
        dist_launcher(run_transformer, mp * pp, config_dict=tiny_config(
            prefix, mp, pp, 1, iters=2, trainer={"save_dir": str(d), "save_interval": 2}))
        out[name] = d
    return out


def test_directory_layout_and_keys(runs):
    d = runs["single"]
    assert (d / "latest").read_text().strip() == "global_step2"
    step = d / "global_step2"
    names = sorted(p.name for p in step.iterdir())
    assert "config.yml" in names and "context_global_rank_0.pt" in names
    model_files = [n for n in names if n.startswith("model_state_layer_")]
    assert model_files == ["model_state_layer_0_EmbeddingInput.pt", "model_state_layer_1_TransformerLayer.pt",
                           "model_state_layer_2_TransformerLayer.pt", "model_state_layer_3_LayerNormWrapper.pt",
                           "model_state_layer_4_TransformerLMHead.pt"]
    assert [n for n in names if n.startswith("optimizer_state_layer_")] == [f"optimizer_state_layer_{i}.pt" for i in range(5)]
    emb = torch.load(step / model_files[0], weights_only=False)
    assert list(emb) == ["embedding.weight"] and emb["embedding.weight"].shape == (128, 64) and emb["embedding.weight"].device.type == "cpu"
    layer = torch.load(step / model_files[1], weights_only=False)
    assert set(layer) == {"input_layernorm.weight", "self_attention.query.weight", "self_attention.key.weight",
                          "self_attention.value.weight", "self_attention.dense.weight", "post_attention_layernorm.weight",
                          "mlp.dense_in.weight", "mlp.siglu_weight.weight", "mlp.dense_out.weight"}
    assert layer["self_attention.key.weight"].shape == (32, 64)  # 2 kv heads x 16
    assert list(torch.load(step / model_files[3], weights_only=False)) == ["norm.weight"]
    assert list(torch.load(step / model_files[4], weights_only=False)) == ["linear.weight"]
    opt = torch.load(step / "optimizer_state_layer_1.pt", weights_only=False)
    assert {"step_index", "loss_scaler", "optimizer_param_groups", "parameters"} <= set(opt)
    key = "layer_index_1_parameter_name_mlp.dense_out.weight_is_model_parallel_True_model_parallel_dimension_1"
    assert key in opt["parameters"], list(opt["parameters"])[:3]
    entry = opt["parameters"][key]
    assert entry["parameter"].dtype == torch.float32 and entry["parameter"].shape == layer["mlp.dense_out.weight"].shape
    assert set(entry["optimizer_state"]) == {"step", "exp_avg", "exp_avg_sq"}
    assert entry["optimizer_state"]["exp_avg"].shape == entry["parameter"].shape
    ctx = torch.load(step / "context_global_rank_0.pt", weights_only=False)
    assert ctx["iterations"] == 2 and ctx["consumed_samples"] == 2 * 2 and "topology" in ctx


@pytest.mark.parametrize("other", ["tp2", "pp2"])
def test_files_are_layout_independent(runs, other):
    a, b = runs["single"] / "global_step2", runs[other] / "global_step2"
    fa = sorted(p.name for p in a.glob("*_state_layer_*.pt"))
    fb = sorted(p.name for p in b.glob("*_state_layer_*.pt"))
    assert fa == fb
    for n in fa:
        if n.startswith("model_state"):
            sa, sb_ = torch.load(a / n, weights_only=False), torch.load(b / n, weights_only=False)
            assert set(sa) == set(sb_) and all(sa[k].shape == sb_[k].shape and sa[k].dtype == sb_[k].dtype for k in sa), n
        else:
            pa, pb = torch.load(a / n, weights_only=False)["parameters"], torch.load(b / n, weights_only=False)["parameters"]
            assert set(pa) == set(pb), n
            assert all(pa[k]["parameter"].shape == pb[k]["parameter"].shape for k in pa)
    world = {"tp2": 2, "pp2": 2}[other]
    assert sorted(p.name for p in b.glob("context_global_rank_*.pt")) == [f"context_global_rank_{r}.pt" for r in range(world)]
