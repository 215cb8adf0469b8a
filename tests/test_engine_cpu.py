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
