"""3-D engine on CPU/gloo with the minimal model: training, layouts, resume, activation checkpointing."""
import pytest

from tests.minimal import run_minimal
from tests.utils import dist_launcher


def base_config(mp=1, pp=1, dp=1, gas=2, zero=False, act="disabled", tying=False, iters=6, **trainer):
    return {
        "topology": {"model_parallel_size": mp, "pipe_parallel_size": pp, "data_parallel_size": dp,
                     "micro_batch_size": 2, "gradient_accumulation_steps": gas,
                     "activation_checkpointing_type": act},
        "optimizer": {"zero": zero, "gradient_clipping": 1.0},
        "trainer": {"train_iterations": iters, "assert_checkpoint_loaded": False, "seed": 42, **trainer},
        "weight_tying": tying,
    }


@pytest.mark.parametrize("mp,pp,dp", [(1, 1, 1), (1, 1, 2), (1, 2, 1), (2, 1, 1), (2, 2, 1), (1, 2, 2)])
@pytest.mark.parametrize("zero", [False, True])
def test_training_runs_and_learns(mp, pp, dp, zero):
    world = mp * pp * dp
    out = dist_launcher(run_minimal, world, config_dict=base_config(mp, pp, dp, zero=zero, iters=8))
    losses = out[0]
    assert len(losses) == 8 and all(l == l for l in losses)
    assert losses[-1] < losses[0]


@pytest.mark.parametrize("mp,pp,dp,zero,tying", [(1, 1, 1, False, False), (1, 2, 1, True, True), (2, 1, 1, True, False), (1, 1, 2, True, False)])
def test_resume_is_exact(tmp_path, mp, pp, dp, zero, tying):
    world = mp * pp * dp
    full = dist_launcher(run_minimal, world, config_dict=base_config(mp, pp, dp, zero=zero, tying=tying, iters=8,
                                                                      save_dir=str(tmp_path), save_interval=5))[0]
    resumed = dist_launcher(run_minimal, world, config_dict=base_config(mp, pp, dp, zero=zero, tying=tying, iters=8,
                                                                         load_dir=str(tmp_path)))[0]
    assert len(resumed) == 3
    assert resumed == full[5:], (full, resumed)


def test_resume_with_other_layout(tmp_path):
    full = dist_launcher(run_minimal, 2, config_dict=base_config(1, 2, 1, zero=True, iters=8, save_dir=str(tmp_path),
                                                                  save_interval=5))[0]
    for mp, pp, dp, zero, gas in [(2, 1, 1, False, 2), (1, 1, 2, True, 1), (1, 1, 1, False, 2)]:
        resumed = dist_launcher(run_minimal, mp * pp * dp, config_dict=base_config(mp, pp, dp, gas=gas, zero=zero, iters=8,
                                                                                    load_dir=str(tmp_path)))[0]
        assert len(resumed) == 3
        for a, b in zip(full[5:], resumed):
            assert abs(a - b) / abs(a) < 0.15, (full, resumed)


@pytest.mark.parametrize("act", ["every_layer", "every_pipe_stage"])
def test_activation_checkpointing_is_exact(act):
    a = dist_launcher(run_minimal, 2, config_dict=base_config(1, 2, 1, iters=4))[0]
    b = dist_launcher(run_minimal, 2, config_dict=base_config(1, 2, 1, act=act, iters=4))[0]
    assert a == b


def test_parameter_count_is_layout_invariant():
    counts = set()
    for mp, pp in [(1, 1), (2, 1), (1, 2), (2, 2)]:
        out = dist_launcher(run_minimal, mp * pp, config_dict=base_config(mp, pp, 1, iters=1, tying=True), return_state=True)
        counts.add(out[0]["params_count"][1])
    assert len(counts) == 1, counts


def _mp_comm_worker(rank: int, world_size: int, master_port: int):
    import torch

    from scaling_b200.core import Topology, TopologyConfig
    from scaling_b200.core.nn import ModelParallelCommunicator

    topo = Topology(TopologyConfig(model_parallel_size=2, pipe_parallel_size=1, data_parallel_size=1, global_rank=rank,
                                   micro_batch_size=1, gradient_accumulation_steps=1))
    topo.initialize_distributed("127.0.0.1", str(master_port))
    comm = ModelParallelCommunicator(topo)
    tree = {"a": torch.arange(6.0).reshape(2, 3), "b": [torch.tensor([True, False]), 7, "x"], "c": None}
    got = comm.sync_data(tree if rank == 0 else None)
    assert got["b"][1:] == [7, "x"] and got["c"] is None
    assert torch.equal(got["a"].cpu(), torch.arange(6.0).reshape(2, 3)) and got["b"][0].tolist() == [True, False]
    # same structure: only payloads travel
    got = comm.sync_data({"a": tree["a"] + 1, "b": tree["b"], "c": None} if rank == 0 else None)
    assert torch.equal(got["a"].cpu(), torch.arange(6.0).reshape(2, 3) + 1)
    # a new structure needs a reset (or continuous recommunication)
    comm.reset_communication_meta(use_continuous_recommunication=True)
    got = comm.sync_data([torch.ones(4, dtype=torch.long)] if rank == 0 else None)
    assert got[0].tolist() == [1, 1, 1, 1]
    got = comm.sync_data((torch.zeros(2), 3) if rank == 0 else None)
    assert got[1] == 3 and got[0].shape == (2,)
    return True


def test_model_parallel_communicator():
    assert all(dist_launcher(_mp_comm_worker, 2).values())


def _eval_worker(rank: int, world_size: int, master_port: int, pp: int, eval_iterations: int = 1):
    """Evaluation runs gradient_accumulation_steps micro batches through the inference schedule under no_grad and
    reports evaluation/* metrics every eval_interval steps; parameters are untouched by it."""
    import torch

    from scaling_b200.core import BaseTrainer, Optimizer, OptimizerParamGroup, OptimizerParamGroupConfig, ParallelModule, Topology
    from tests.minimal import (
        MinimalConfig,
        MinimalContext,
        MinimalDataset,
        layer_specs,
        loss_function,
        metrics_aggregation_fn,
    )

    cfg = MinimalConfig.from_dict({
        "topology": {"model_parallel_size": 1, "pipe_parallel_size": pp, "data_parallel_size": 1, "micro_batch_size": 2,
                     "gradient_accumulation_steps": 2, "world_size": world_size, "global_rank": rank, "local_slot": rank},
        "trainer": {"train_iterations": 4, "assert_checkpoint_loaded": False, "eval_interval": 2, "seed": 1,
                    "dataloader_pin_memory": False, "eval_iterations": eval_iterations}})
    topo = Topology(cfg.topology)
    ctx = MinimalContext(config=cfg, topology=topo)
    ctx.initialize(master_addr="127.0.0.1", master_port=str(master_port), seed=1)
    model = ParallelModule(layer_specs(cfg, topo), topology=topo, profiler_config=cfg.profiler)
    named = model.named_parameters_with_meta()
    opt = Optimizer(cfg.optimizer, [OptimizerParamGroup(named, OptimizerParamGroupConfig(
        name="all", weight_decay=0.0, learning_rate_scheduler=cfg.learning_rate_scheduler))], topo)
    data = MinimalDataset(seed=1) if topo.is_io_rank else None
    evald = MinimalDataset(seed=2, n=64) if topo.is_io_rank else None
    trainer = BaseTrainer(config=cfg.trainer, context=ctx, parallel_module=model, optimizer=opt, dataset=data,
                          dataset_evaluation=evald, sync_batch_to_model_parallel=MinimalDataset.sync_batch_to_model_parallel,
                          loss_function=loss_function, metrics_aggregation_fn=metrics_aggregation_fn)
    before = [p.detach().clone() for p in model.parameters()]
    out = trainer.eval_step()
    assert all(torch.equal(a, b) for a, b in zip(before, model.parameters()))
    assert all(p.grad is None or float(p.grad.abs().sum()) == 0.0 for p in model.parameters())
    metrics = trainer.run_training(return_metrics=True)
    return {"eval_loss": out.loss, "keys": [sorted(k for k in m if k.startswith("evaluation/")) for m in (metrics or [])],
            "consumed_eval": ctx.consumed_samples_evaluation}


@pytest.mark.parametrize("pp", [1, 2])
def test_evaluation_step(pp):
    res = dist_launcher(_eval_worker, pp, pp=pp)[0]
    assert res["eval_loss"] > 0
    assert [bool(k) for k in res["keys"]] == [False, True, False, True]
    assert "evaluation/loss" in res["keys"][1]
    assert res["consumed_eval"] == 3 * 4  # one explicit + two scheduled evaluations of global batch 4


def test_eval_iterations_average_several_steps():
    res = dist_launcher(_eval_worker, 1, pp=1, eval_iterations=3)[0]
    assert res["consumed_eval"] == (1 + 2 * 3) * 4  # one explicit step + two evaluations of three steps each
    assert "evaluation/loss" in res["keys"][1]


def test_non_distributed_mode_and_buffers_in_checkpoint(tmp_path):
    """`context.initialize(distributed=False)` only sets the device; persistent buffers are part of the layer files."""
    import torch

    from scaling_b200.core import ParallelModule, Topology
    from tests.minimal import Batch, MinimalConfig, MinimalContext, layer_specs

    cfg = MinimalConfig.from_dict({"topology": {"model_parallel_size": 1, "pipe_parallel_size": 1, "data_parallel_size": 1,
                                                "micro_batch_size": 2, "gradient_accumulation_steps": 1, "world_size": 1,
                                                "global_rank": 0, "local_slot": 0}})
    topo = Topology(cfg.topology)
    ctx = MinimalContext(config=cfg, topology=topo)
    ctx.initialize(master_addr="127.0.0.1", master_port="0", seed=3, distributed=False)
    assert not torch.distributed.is_initialized() and not topo.is_distributed_initialized
    model = ParallelModule(layer_specs(cfg, topo), topology=topo, profiler_config=cfg.profiler)
    x = Batch(inputs=torch.randint(0, 32, (2, 8)))
    for _, layer in model.local_layers():
        x = layer(x)
    assert x.activations.shape == (2, 8, 32)
    for _, layer in model.local_layers():
        if hasattr(layer, "buffer_for_testing"):
            layer.buffer_for_testing.fill_(7.0)
    model.save_checkpoint(tmp_path)
    files = sorted(p.name for p in tmp_path.glob("model_state_layer_*.pt"))
    assert files == ["model_state_layer_0_EmbeddingLayer.pt", "model_state_layer_1_HiddenLayer.pt",
                     "model_state_layer_2_HiddenLayer.pt", "model_state_layer_3_HeadLayer.pt"]
    sd = torch.load(tmp_path / files[1], weights_only=False)
    assert "buffer_for_testing" in sd and float(sd["buffer_for_testing"][0]) == 7.0
    fresh = ParallelModule(layer_specs(cfg, topo), topology=topo, profiler_config=cfg.profiler)
    fresh.load_checkpoint(tmp_path)
    for (_, a), (_, b) in zip(model.local_layers(), fresh.local_layers()):
        for (n1, t1), (n2, t2) in zip(a.state_dict().items(), b.state_dict().items()):
            assert n1 == n2 and torch.equal(t1, t2), n1


def test_profiler_output_feeds_the_schedule_simulator(tmp_path):
    """Instruction profiler → profile.json (rank 0 gathers every rank's observations) → SimulationEngine replay."""
    import json

    from scaling_b200.core.nn import PipelineScheduleTrain, SimulationEngine

    out = tmp_path / "profile.json"
    cfg = {"topology": {"model_parallel_size": 1, "pipe_parallel_size": 2, "data_parallel_size": 1, "micro_batch_size": 2,
                        "gradient_accumulation_steps": 4},
           "trainer": {"train_iterations": 6, "assert_checkpoint_loaded": False, "seed": 1, "dataloader_pin_memory": False},
           "profiler": {"profile_steps": 3, "profile_start_at_step": 2, "profiler_output": str(out)}}
    dist_launcher(run_minimal, 2, config_dict=cfg)
    data = json.loads(out.read_text())
    assert data["pipe_parallel_size"] == 2 and data["gradient_accumulation_steps"] == 4
    obs = data["observations"]
    assert {o["pipe_parallel_rank"] for o in obs} == {0, 1}
    names = {o["timer_name"] for o in obs}
    assert {"ForwardPass", "BackwardPass", "SendActivation", "RecvActivation", "SendGrad", "RecvGrad", "OptimizerStep"} <= names
    assert {o["step"] for o in obs} == {2, 3, 4} and all(o["duration"] >= 0 for o in obs)
    sim = SimulationEngine.from_profile(PipelineScheduleTrain, out)
    sim.simulate()
    summary = sim.summarize()
    assert summary["duration"] > 0 and len(summary["stages"]) == 2 and 0 <= summary["idle_percent"] < 100
    assert sim.visualize() is not None
    assert PipelineScheduleTrain.visualize_profile(out) is not None


def _missing_entry_worker(rank, world_size, master_port, save_dir):
    """Loads a checkpoint whose optimizer files lack one parameter and steps once: that parameter must continue from
    its LOADED weight (master refreshed from the model files), not from the fresh construction-time master."""
    import torch

    from scaling_b200.core import BaseTrainer, Optimizer, OptimizerParamGroup, OptimizerParamGroupConfig, ParallelModule, Topology
    from tests.minimal import (MinimalConfig, MinimalContext, MinimalDataset, layer_specs, loss_function,
                               metrics_aggregation_fn)

    cfg = base_config(1, 1, 1, zero=True, iters=3, load_dir=save_dir)
    cfg["topology"] = dict(cfg["topology"], world_size=world_size, global_rank=rank, local_slot=rank)
    cfg["trainer"]["seed"] = 7  # a different seed: construction-time weights differ from the saved ones
    config = MinimalConfig.from_dict(cfg)
    topology = Topology(config.topology)
    context = MinimalContext(config=config, topology=topology)
    context.initialize(master_addr="127.0.0.1", master_port=str(master_port), seed=config.trainer.seed)
    model = ParallelModule(layer_specs(config, topology), topology=topology, profiler_config=config.profiler)
    named = model.named_parameters_with_meta()
    group = OptimizerParamGroup(named, OptimizerParamGroupConfig(name="all", weight_decay=0.0,
                                                                 learning_rate_scheduler=config.learning_rate_scheduler))
    optimizer = Optimizer(config.optimizer, [group], topology)
    trainer = BaseTrainer(config=config.trainer, context=context, parallel_module=model, optimizer=optimizer,
                          dataset=MinimalDataset(seed=1), sync_batch_to_model_parallel=MinimalDataset.sync_batch_to_model_parallel,
                          loss_function=loss_function, metrics_aggregation_fn=metrics_aggregation_fn)
    victim = next(p for n, p in model.named_parameters() if n.endswith("down.weight"))
    saved = torch.load(f"{save_dir}/global_step2/model_state_layer_1_HiddenLayer.pt", weights_only=False)["down.weight"]
    assert torch.equal(victim.detach().cpu(), saved)
    trainer.train_step()
    # one AdamW step at lr 0.01 moves every element by at most ~lr; the fresh random master would be far away
    return float((victim.detach().cpu() - saved).abs().max())


def test_optimizer_load_keeps_loaded_weight_for_parameter_without_state(tmp_path):
    import torch

    dist_launcher(run_minimal, 1, config_dict=base_config(1, 1, 1, zero=True, iters=2, save_dir=str(tmp_path), save_interval=2))
    f = tmp_path / "global_step2" / "optimizer_state_layer_1.pt"
    sd = torch.load(f, weights_only=False)
    dropped = [k for k in sd["parameters"] if "down.weight" in k]
    assert len(dropped) == 1
    del sd["parameters"][dropped[0]]
    torch.save(sd, f)
    drift = dist_launcher(_missing_entry_worker, 1, save_dir=str(tmp_path))[0]
    assert drift < 0.05, drift
