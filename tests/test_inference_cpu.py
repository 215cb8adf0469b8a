"""Inference module: load a TP=2 training checkpoint into a single process, logits / recorder / generation, samplers
(reference oracles: tests/transformer/test_inference.py)."""
from pathlib import Path

import pytest
import torch

from tests.transformer_utils import run_transformer, tiny_config, write_dataset
from tests.utils import dist_launcher


def _vocab_file(path: Path, size: int = 128) -> Path:
    from tokenizers import Tokenizer as HF
    from tokenizers import models, pre_tokenizers

    vocab = {"<|endoftext|>": 0}
    for i in range(1, size):
        vocab[f"t{i}"] = i
    tok = HF(models.WordLevel(vocab, unk_token="t1"))
    tok.pre_tokenizer = pre_tokenizers.Whitespace()
    tok.save(str(path))
    return path


@pytest.fixture(scope="module")
def checkpoint(tmp_path_factory):
    root = tmp_path_factory.mktemp("inf")
    prefix = root / "data" / "tokens"
    write_dataset(prefix)
    ckpt = root / "ckpt"
    cfg = tiny_config(prefix, 2, 1, 1, iters=30, trainer={"save_dir": str(ckpt), "save_interval": 30})
    dist_launcher(run_transformer, 2, config_dict=cfg)
    step_dir = ckpt / "global_step30"
    assert step_dir.is_dir()
    _vocab_file(step_dir / "vocab.json")
    return step_dir


def test_logits_generate_and_recorder(checkpoint):
    from scaling_b200.core.nn import RecorderSetting
    from scaling_b200.transformer.inference import TransformerInferenceModule, sample_argmax

    model = TransformerInferenceModule.from_checkpoint(checkpoint, devices=["cpu"])
    prompt = [5, 16, 49, 21, 64]
    logits = model.logits(input_tokens=prompt)
    assert logits.shape == (len(prompt), 128)
    # the training data follows next = (3 t + 1) mod 127 + 1; 30 steps are enough to beat chance by a wide margin
    two = TransformerInferenceModule.from_checkpoint(checkpoint, devices=["cpu", "cpu"])
    torch.testing.assert_close(two.logits(input_tokens=prompt), logits)

    cached = model.generate(max_tokens=6, input_tokens=prompt, stop_tokens=[0], use_cache=True)
    uncached = model.generate(max_tokens=6, input_tokens=prompt, stop_tokens=[0], use_cache=False)
    assert cached.completion_tokens == uncached.completion_tokens
    assert len(cached.completion_tokens) <= 6 and cached.completion_logits.shape[-1] == 128
    torch.testing.assert_close(cached.completion_logits[0], uncached.completion_logits[0], rtol=1e-4, atol=1e-4)
    assert cached.completion_text is not None
    assert model.generate(max_tokens=3, input_text="t5 t16 t49", stop_tokens=[0], sample_fn=sample_argmax).completion_tokens

    out, record = model.logits_with_hidden_state_recorder(
        input_tokens=prompt, recorder_settings_per_layer={1: RecorderSetting(), 2: RecorderSetting(include_modules=["self_attention", "mlp"])})
    torch.testing.assert_close(out, logits)
    assert set(record) == {1, 2} and "" in record[1] and set(record[2]) == {"self_attention", "mlp"}
    assert record[1][""].activations.shape[-1] == 64


def test_samplers():
    from scaling_b200.transformer.inference import (
        sample_argmax,
        sample_temperature,
        sample_top_k,
        sample_top_p,
        top_k_transform,
        top_p_transform,
    )

    torch.manual_seed(0)
    logits = torch.randn(2, 3, 50)
    assert sample_argmax(logits).tolist() == logits[:, -1].argmax(-1).tolist()
    k = top_k_transform(logits, 5)
    assert (torch.isfinite(k).sum(-1) == 5).all()
    p = top_p_transform(logits, 0.5)
    kept = torch.softmax(logits, -1).masked_fill(~torch.isfinite(p), 0).sum(-1)
    assert (kept >= 0.5).all() and (torch.isfinite(p).sum(-1) >= 1).all()
    # dropping the last kept token must fall below the threshold (minimal nucleus)
    probs = torch.softmax(logits, -1)
    smallest_kept = probs.masked_fill(~torch.isfinite(p), 2.0).min(-1).values
    assert ((kept - smallest_kept) <= 0.5 + 1e-6).all()
    for fn in (sample_temperature, lambda x: sample_top_k(x, 3), lambda x: sample_top_p(x, 0.3)):
        s = fn(logits)
        assert s.shape == (2,) and ((0 <= s) & (s < 50)).all()
    peaked = torch.full((1, 1, 10), -50.0)
    peaked[0, 0, 7] = 50.0
    assert sample_temperature(peaked, 0.7).item() == 7 and sample_top_p(peaked).item() == 7
