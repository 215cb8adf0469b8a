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
