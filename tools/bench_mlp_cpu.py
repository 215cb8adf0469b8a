"""BASELINE config 1: the MLP example's 3-D-parallel plumbing on CPU / gloo, world 2 (the reference cannot run this
configuration at all: it requires CUDA + NCCL).  Prints steps/s of `BaseTrainer.train_step` for a few layouts."""
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from tests.utils import dist_launcher  # noqa: E402


def worker(rank: int, world_size: int, master_port: int, layout, steps: int):
    import yaml

    from examples.mlp_example import model as M
    from examples.mlp_example.config import MLPConfig
    from examples.mlp_example.data import MNISTDataset
    from scaling_b200.core import BaseTrainer, LaunchConfig, Topology
    from scaling_b200.core.config import overwrite_recursive

    mp, pp, dp = layout
    payload = yaml.load(open(ROOT / "examples/mlp_example/config.yml"), Loader=yaml.FullLoader)
    overwrite_recursive(payload, {"topology": {"model_parallel_size": mp, "pipe_parallel_size": pp, "data_parallel_size": dp,
                                               "micro_batch_size": 256 // dp, "gradient_accumulation_steps": 1,
                                               "global_batch_size": None, "world_size": None},
                                  "architecture": {"hidden_dim": 128, "n_hidden_layers": 1},
                                  "logger": {"log_level": "error", "use_tensorboard": False, "log_dir": None},
                                  "data": {"root": "/nonexistent", "synthetic_size": 8192}})
    launch = LaunchConfig(master_port=master_port, master_addr="127.0.0.1", world_size=world_size, global_rank=rank,
                          local_slot=rank, payload=payload)
    config = MLPConfig.from_dict(launch.overwrite_config_dict_with_launcher_args(dict(payload)))
    topology = Topology(config=config.topology)
    context = M.MLPContext(config=config, topology=topology)
    context.initialize(master_addr="127.0.0.1", master_port=str(master_port), seed=1)
    module = M.init_model(context=context)
    data = MNISTDataset(root="/nonexistent", train=True, synthetic_size=8192) if topology.is_io_rank else None
    trainer = BaseTrainer(config=config.trainer, context=context, parallel_module=module,
                          optimizer=M.init_optimizer(context=context, model=module), dataset=data,
                          sync_batch_to_model_parallel=MNISTDataset.sync_batch_to_model_parallel,
                          loss_function=M.loss_function, metrics_aggregation_fn=M.metrics_aggregation_fn)
    for _ in range(5):
        trainer.train_step()
    t0 = time.perf_counter()
    for _ in range(steps):
        out = trainer.train_step()
    return {"steps_per_s": steps / (time.perf_counter() - t0), "loss": out.loss}


if __name__ == "__main__":
    res = {}
    for name, layout in {"tp2": (2, 1, 1), "pp2": (1, 2, 1), "dp2": (1, 1, 2)}.items():
        r = dist_launcher(worker, 2, layout=layout, steps=100)
        res[name] = {"steps_per_s": min(v["steps_per_s"] for v in r.values()), "final_loss": r[0]["loss"]}
        print(name, res[name], flush=True)
    Path("profiles").mkdir(exist_ok=True)
    json.dump({"config": "MLP example 784-128-128-128-10, global batch 256, fp32, CPU/gloo, world 2, synthetic digits",
               "metric": "optimizer steps/s (min over ranks)", "results": res,
               "reference": "unavailable: requires CUDA + NCCL (topology.py:109-117)"},
              open("profiles/bench_r1_mlp_cpu.json", "w"), indent=1)
