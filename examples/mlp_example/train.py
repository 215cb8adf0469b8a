"""Per-process entry point of the MLP example: the launcher starts this file once per device slot."""
from scaling_b200.core import BaseTrainer, LaunchConfig, Topology
from scaling_b200.core.logging import logger

from examples.mlp_example import model as M
from examples.mlp_example.config import MLPConfig
from examples.mlp_example.data import MNISTDataset


def main(launch_config: LaunchConfig, return_metrics: bool = False):
    payload = launch_config.overwrite_config_dict_with_launcher_args(dict(launch_config.payload or {}))
    config = MLPConfig.from_dict(payload)
    topology = Topology(config=config.topology)
    rank = topology.config.global_rank
    logger.configure(config=config.logger, name=f"RANK {rank}", global_rank=rank)
    context = M.MLPContext(config=config, topology=topology)
    context.initialize(master_addr=launch_config.master_addr, master_port=str(launch_config.master_port),
                       seed=config.trainer.seed)
    module = M.init_model(context=context)
    splits = {}
    if topology.is_io_rank:  # only the first / last pipeline stage of TP rank 0 read data
        for name, is_train in (("dataset", True), ("dataset_evaluation", False)):
            splits[name] = MNISTDataset(root=config.data.root, train=is_train, synthetic_size=config.data.synthetic_size)
    trainer = BaseTrainer(config=config.trainer, context=context, parallel_module=module,
                          optimizer=M.init_optimizer(context=context, model=module),
                          sync_batch_to_model_parallel=MNISTDataset.sync_batch_to_model_parallel,
                          loss_function=M.loss_function, metrics_aggregation_fn=M.metrics_aggregation_fn,
                          dataset=splits.get("dataset"), dataset_evaluation=splits.get("dataset_evaluation"))
    return trainer.run_training(return_metrics=return_metrics)


if __name__ == "__main__":
    main(LaunchConfig.from_launcher_args())
