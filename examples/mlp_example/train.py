"""Per-process entry point of the MLP example (started once per device by the launcher)."""
from scaling_b200.core import BaseTrainer
from scaling_b200.core.logging import logger
from scaling_b200.core.runner import LaunchConfig
from scaling_b200.core.topology import Topology

from examples.mlp_example.config import MLPConfig
from examples.mlp_example.context import MLPContext
from examples.mlp_example.data import MNISTDataset
from examples.mlp_example.model import init_model, init_optimizer, loss_function, metrics_aggregation_fn


def main(launch_config: LaunchConfig, return_metrics: bool = False):
    payload = dict(launch_config.payload or {})
    payload["topology"] = dict(payload.get("topology", {}), world_size=launch_config.world_size,
                               global_rank=launch_config.global_rank, local_slot=launch_config.local_slot)
    config = MLPConfig.from_dict(payload)
    topology = Topology(config=config.topology)
    context = MLPContext(config=config, topology=topology)
    logger.configure(config=config.logger, name=f"RANK {topology.config.global_rank}", global_rank=topology.config.global_rank)
    context.initialize(master_addr=launch_config.master_addr, master_port=str(launch_config.master_port),
                       seed=config.trainer.seed)
    model = init_model(context=context)
    optimizer = init_optimizer(context=context, model=model)
    train_data = valid_data = None
    if topology.is_io_rank:
        train_data = MNISTDataset(root=config.data.root, train=True, synthetic_size=config.data.synthetic_size)
        valid_data = MNISTDataset(root=config.data.root, train=False, synthetic_size=config.data.synthetic_size)
    trainer = BaseTrainer(config=context.config.trainer, context=context, parallel_module=model, optimizer=optimizer,
                          dataset=train_data, dataset_evaluation=valid_data,
                          sync_batch_to_model_parallel=MNISTDataset.sync_batch_to_model_parallel,
                          metrics_aggregation_fn=metrics_aggregation_fn, loss_function=loss_function)
    return trainer.run_training(return_metrics=return_metrics)


if __name__ == "__main__":
    main(LaunchConfig.from_launcher_args())
