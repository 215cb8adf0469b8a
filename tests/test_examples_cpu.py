"""The shipped examples run: MLP example trains on its synthetic digits with TP=2 (gloo), example configs parse."""
from pathlib import Path

from tests.utils import dist_launcher

ROOT = Path(__file__).resolve().parents[1]


def _mlp_worker(rank: int, world_size: int, master_port: int, overrides: dict):
    import yaml

    from examples.mlp_example.train import main
    from scaling_b200.core import LaunchConfig
    from scaling_b200.core.config import overwrite_recursive

    payload = yaml.load(open(ROOT / "examples/mlp_example/config.yml"), Loader=yaml.FullLoader)
    overwrite_recursive(payload, overrides)
    launch = LaunchConfig(master_port=master_port, master_addr="127.0.0.1", world_size=world_size, global_rank=rank,
                          local_slot=rank, payload=payload)
    metrics = main(launch, return_metrics=True)
    return [(m["training/loss"], m.get("training/accuracy")) for m in (metrics or [])]


def test_mlp_example_trains(tmp_path):
    over = {"trainer": {"train_iterations": 60}, "topology": {"micro_batch_size": 32, "global_batch_size": None},
            "data": {"root": str(tmp_path / "no_mnist_here"), "synthetic_size": 2048},
            "logger": {"log_level": "warning", "use_tensorboard": False, "log_dir": None}}
    out = dist_launcher(_mlp_worker, 2, overrides=over)[0]
    assert len(out) == 60
    assert out[-1][0] < 0.5 * out[0][0], out
    assert out[-1][1] is None or float(out[-1][1]) > 0.7


def test_example_configs_parse():
    from examples.mlp_example.config import MLPConfig
    from scaling_b200.transformer import TransformerConfig

    m = MLPConfig.from_yaml(ROOT / "examples/mlp_example/config.yml")
    assert m.topology.world_size == 2 and m.runner.script is not None
    t = TransformerConfig.from_yaml(ROOT / "examples/transformer_example/config.yml")
    assert t.topology.world_size == 8 and t.transformer_architecture.attention_num_kv_heads == 4
    # the annotated rendering is a faithful round trip
    assert TransformerConfig.from_dict(__import__("yaml").load(t.as_annotated_str(), Loader=__import__("yaml").FullLoader)) == t
