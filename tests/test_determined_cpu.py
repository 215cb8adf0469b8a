"""Determined glue without Determined: hyper-parameter → config mapping and the no-op fallbacks."""
from scaling_b200.core.determined import init, maybe_periodic_stacktraces
from scaling_b200.transformer.train_determined import hparams_to_overwrite


def test_hparam_mapping():
    hp = {"learning_rate": 1e-3, "target_train_tokens": 2 ** 20, "warmup_tokens": 2 ** 16, "sequence_length": 1024,
          "global_batch_size": 32, "model_parallel_size": 2, "kernel": "flash_attention", "wandb_project": "p"}
    o = hparams_to_overwrite(hp, {"runner": {"use_determined": True}})
    assert o["trainer"]["train_iterations"] == 32 and o["learning_rate_scheduler"]["learning_rate_warmup_steps"] == 2
    assert o["learning_rate_scheduler"]["learning_rate"] == 1e-3 and o["topology"]["model_parallel_size"] == 2
    assert o["transformer_architecture"]["masked_softmax"] == {"kernel": "flash_attention"}
    assert o["transformer_architecture"]["sequence_length"] == 1024 and o["logger"]["wandb_project"] == "p"
    assert o["runner"] == {"use_determined": True}
    # explicit iterations win; a restart layout replaces everything else
    assert hparams_to_overwrite({**hp, "train_iterations": 7})["trainer"]["train_iterations"] == 7
    assert hparams_to_overwrite({"learning_rate": 1.0, "layout": {"micro_batch_size": 3}})["topology"] == {"micro_batch_size": 3}
    assert hparams_to_overwrite(None, {"a": 1}) == {"a": 1} and hparams_to_overwrite({}) == {}


def test_noop_fallbacks():
    with maybe_periodic_stacktraces(False):
        pass
    with maybe_periodic_stacktraces(True, interval_seconds=3600):
        pass
    with init() as ctx:
        assert ctx is None
