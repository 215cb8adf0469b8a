"""Transformer suite end to end on CPU/gloo: training, TP/PP/DP layouts, sequence parallel, resume, finetuning."""
import pytest

from tests.transformer_utils import run_transformer, tiny_config, write_dataset
from tests.utils import dist_launcher


@pytest.fixture()
def data(tmp_path):
    prefix = tmp_path / "data" / "tokens"
    write_dataset(prefix)
    return prefix


@pytest.mark.parametrize("mp,pp,dp", [(1, 1, 1), (2, 1, 1), (1, 2, 1), (1, 1, 2), (2, 2, 1)])
def test_transformer_trains(data, mp, pp, dp):
    losses = dist_launcher(run_transformer, mp * pp * dp, config_dict=tiny_config(data, mp, pp, dp, gas=2, iters=12))[0]
    assert len(losses) == 12
    assert losses[-1] < losses[0] * 0.9, losses


def test_sequence_parallel_matches_tensor_parallel(tmp_path, data):
    """SP only re-shards norm/residual regions; starting from the same checkpoint its loss curve must follow plain TP."""
    ckpt = tmp_path / "ckpt"
    dist_launcher(run_transformer, 2, config_dict=tiny_config(data, 2, 1, 1, iters=1, trainer={"save_dir": str(ckpt), "save_interval": 1}))
    tp = dist_launcher(run_transformer, 2, config_dict=tiny_config(data, 2, 1, 1, iters=5, trainer={"load_dir": str(ckpt)}))[0]
    sp = dist_launcher(run_transformer, 2, config_dict=tiny_config(data, 2, 1, 1, iters=5, trainer={"load_dir": str(ckpt)},
                                                                   topology={"sequence_parallel": True}))[0]
    assert len(tp) == len(sp) == 4
    for a, b in zip(tp, sp):
        assert abs(a - b) / abs(a) < 1e-3, (tp, sp)


@pytest.mark.parametrize("mp,pp", [(1, 1), (2, 1), (1, 2)])
def test_transformer_resume_exact(tmp_path, data, mp, pp):
    ckpt = tmp_path / "ckpt"
    world = mp * pp
    # every kind of dropout: the counter-based ones (embedding, after attention / MLP) and the attention-probability
    # dropout of the dense path, which samples torch's default generator (reference acceptance test:
    # tests/transformer/test_training.py:150-300 resumes at 1e-10 with all four at 0.1)
    drop = {"dropout_after_mlp": 0.1, "dropout_after_attention": 0.1, "dropout_embedding": 0.1, "dropout_attention_probs": 0.1}
    full = dist_launcher(run_transformer, world, config_dict=tiny_config(
        data, mp, pp, 1, iters=8, trainer={"save_dir": str(ckpt), "save_interval": 5}, arch=drop))[0]
    resumed = dist_launcher(run_transformer, world, config_dict=tiny_config(
        data, mp, pp, 1, iters=8, trainer={"load_dir": str(ckpt)}, arch=drop))[0]
    assert resumed == full[5:], (full, resumed)


def test_weight_tying_and_layernorm_bias_model(data):
    cfg = tiny_config(data, 1, 2, 1, gas=2, iters=10, arch={
        "weight_tying": True, "norm_type": "layernorm", "mlp_type": "default", "mlp_factor": 4.0, "attention_bias": True,
        "mlp_bias": True, "attention_qkv_in_one": True, "attention_num_kv_heads": None,
        "relative_position_embedding_type": "rotary"})
    losses = dist_launcher(run_transformer, 2, config_dict=cfg)[0]
    assert losses[-1] < losses[0]


@pytest.mark.parametrize("arch", [
    {"attention_num_kv_heads": 1},  # true MQA: the single KV head is replicated on both TP ranks
    {"attention_num_kv_heads": 2, "key_query_norm": True},  # per-head QK-norm weights are TP-replicated
])
def test_tp_replicated_parameters_follow_single_gpu_run(tmp_path, data, arch):
    """Replicated-but-partially-used parameters (MQA K/V projections, QK-norm weights) need their gradients summed
    over the TP ranks that share them; with that, TP=2 follows the TP=1 loss curve from the same checkpoint."""
    ckpt = tmp_path / "ckpt"
    dist_launcher(run_transformer, 1, config_dict=tiny_config(data, 1, 1, 1, iters=1, arch=arch,
                                                              trainer={"save_dir": str(ckpt), "save_interval": 1}))
    one = dist_launcher(run_transformer, 1, config_dict=tiny_config(data, 1, 1, 1, iters=6, arch=arch,
                                                                    trainer={"load_dir": str(ckpt)}))[0]
    two = dist_launcher(run_transformer, 2, config_dict=tiny_config(data, 2, 1, 1, iters=6, arch=arch,
                                                                    trainer={"load_dir": str(ckpt)}))[0]
    assert len(one) == len(two) == 5
    for a, b in zip(one, two):
        assert abs(a - b) / abs(a) < 2e-3, (one, two)
