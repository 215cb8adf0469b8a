"""Finetuning datasets: prompt/completion (json + memory map), chat jsonl, images, soft prompt, TP broadcast
(reference oracles: tests/transformer/test_data/test_finetuning_*)."""
import json
from pathlib import Path

import numpy as np
import pytest
import torch

from scaling_b200.transformer.data.finetuning import (
    IMAGE_ENCODER_TOKEN_COUNTS,
    FinetuningChatDataset,
    FinetuningTextDataset,
)
from scaling_b200.transformer.tokenizer import load_tokenizers

WORDS = "the quick brown fox jumps over lazy dog hello world a b c d e f g yes no maybe image says".split()


@pytest.fixture(scope="module")
def toks(tmp_path_factory):
    from tokenizers import Tokenizer as HF
    from tokenizers import decoders, models, pre_tokenizers

    d = tmp_path_factory.mktemp("tok")
    vocab = {"<|endoftext|>": 0, "<unk>": 1}
    for w in WORDS:
        vocab["▁" + w] = len(vocab)
        vocab[w] = len(vocab)
    tok = HF(models.WordLevel(vocab, unk_token="<unk>"))
    tok.pre_tokenizer = pre_tokenizers.Metaspace(replacement="▁", prepend_scheme="always")
    tok.decoder = decoders.Metaspace(replacement="▁")
    tok.save(str(d / "vocab.json"))
    return load_tokenizers(d / "vocab.json")


def test_prompt_completion_json(tmp_path, toks):
    tok, tok_nps = toks
    recs = [{"prompt": "the quick brown fox", "completion": "jumps over"}, {"prompt": "hello", "completion": "world"}]
    path = tmp_path / "data.json"
    path.write_text(json.dumps(recs))
    ds = FinetuningTextDataset(path, sequence_length=12, seed=1, softprompt_n_tokens=0, tokenizer=tok,
                               tokenizer_no_prefix_space=tok_nps, shuffle=False)
    assert len(ds) == 2
    item = ds[0]
    p, c = tok.encode(recs[0]["prompt"]), tok.encode(recs[0]["completion"])
    full = p + c
    assert item.input_token_ids.tolist()[:len(full)] == full
    assert item.input_token_ids.shape == (12,) and item.target_token_ids.shape == (12,)
    assert item.target_token_ids.tolist()[:len(full) - 1] == full[1:]
    # loss exactly on the completion tokens and on ONE end-of-text after them
    w = item.loss_weights.tolist()
    expect = [0.0] * (len(p) - 1) + [1.0] * (len(c) + 1)
    assert w == expect + [0.0] * (12 - len(expect))
    assert item.target_token_ids[len(full) - 1].item() == tok.eos_token_id
    batch = ds.collate([ds[0], ds[1]])
    assert batch.input_token_ids.shape == (2, 12)
    assert batch.cumulative_seq_lengths.tolist() == [0, 12, 24]
    assert batch.input_images is None


def test_prompt_completion_memory_map_and_softprompt(tmp_path, toks):
    tok, tok_nps = toks
    jl = tmp_path / "d.jsonl"
    jl.write_text("\n".join(json.dumps(r) for r in [{"prompt": "a b c", "completion": "d e"}, {"prompt": "f", "completion": "g yes no"}]))
    FinetuningTextDataset.convert_jsonl(jl, tok, tok_nps, tmp_path / "mm")
    ds = FinetuningTextDataset(tmp_path / "mm", sequence_length=10, seed=3, softprompt_n_tokens=2, tokenizer=tok,
                               tokenizer_no_prefix_space=tok_nps, memory_map_dataset=True, shuffle=False)
    item = ds[0]
    p, c = tok.encode("a b c"), tok_nps.encode("d e")
    assert item.input_token_ids.tolist()[:2] == [0, 0]
    assert item.input_token_ids.tolist()[2:2 + len(p) + len(c)] == p + c
    w = item.loss_weights.tolist()
    assert w[:2 + len(p) - 1] == [0.0] * (2 + len(p) - 1)
    assert sum(w) == len(c) + 1
    # too long: truncated to seq+1 tokens, no padding target
    ds_short = FinetuningTextDataset(tmp_path / "mm", sequence_length=3, seed=3, softprompt_n_tokens=0, tokenizer=tok,
                                     tokenizer_no_prefix_space=tok_nps, memory_map_dataset=True, shuffle=False)
    it = ds_short[0]
    assert it.input_token_ids.tolist() == (p + c)[:3] and it.target_token_ids.tolist() == (p + c)[1:4]
    # shuffling is a deterministic function of the seed
    a = FinetuningTextDataset(tmp_path / "mm", 10, 7, 0, tok, tok_nps, memory_map_dataset=True).data
    b = FinetuningTextDataset(tmp_path / "mm", 10, 7, 0, tok, tok_nps, memory_map_dataset=True).data
    assert a == b


def test_chat_with_image(tmp_path, toks):
    from PIL import Image

    tok, tok_nps = toks
    Image.fromarray((np.random.rand(40, 60, 3) * 255).astype("uint8")).save(tmp_path / "pic.jpg")
    conv = [{"type": "text", "content": "hello"}, {"type": "image", "content": "pic.jpg"},
            {"type": "text", "content": "image says", "has_loss": False}, {"type": "text", "content": "yes", "has_loss": True}]
    path = tmp_path / "chat.jsonl"
    path.write_text(json.dumps(conv) + "\n" + json.dumps([{"type": "text", "content": "no maybe", "has_loss": True}]) + "\n")
    L = IMAGE_ENCODER_TOKEN_COUNTS + 16
    ds = FinetuningChatDataset(path, sequence_length=L, seed=0, softprompt_n_tokens=0, tokenizer=tok,
                               tokenizer_no_prefix_space=tok_nps, shuffle=False)
    item = ds[0]
    n_hello = len(tok.encode("hello"))
    assert item.input_image_locations == [(n_hello, n_hello + IMAGE_ENCODER_TOKEN_COUNTS)]
    assert item.input_images[0].shape == (3, 384, 384)
    # only the last turn is a target; weight i belongs to target token i+1
    n_before = n_hello + IMAGE_ENCODER_TOKEN_COUNTS + len(tok_nps.encode("image says"))
    n_yes = len(tok_nps.encode("yes"))
    w = item.loss_weights
    assert w.sum().item() == n_yes and w[n_before - 1:n_before - 1 + n_yes].sum().item() == n_yes
    batch = ds.collate([ds[0], ds[1]])
    assert batch.input_images.shape == (1, 3, 384, 384)
    assert batch.input_image_locations.tolist() == [[0, n_hello, n_hello + IMAGE_ENCODER_TOKEN_COUNTS]]
    synced = FinetuningChatDataset.sync_batch_to_model_parallel(None, batch)
    assert synced is batch


def test_image_encoder_shapes():
    from scaling_b200.core.nn.norms import LayerNormConfig
    from scaling_b200.transformer.model.image_encoder import ClipModifiedResNet, ImageEncoder

    trunk = ClipModifiedResNet([1, 1, 1, 1], 16)
    assert trunk(torch.randn(2, 3, 64, 64)).shape == (2, 4, 512)
    enc = ImageEncoder(32, torch.device("cpu"), layernorm_config=LayerNormConfig())
    names = dict(enc.named_parameters())
    assert "input_encoder.layer3.17.conv2.weight" in names and "proj.weight" in names and "layernorm.weight" in names
    assert names["proj.weight"].shape == (32, 3072)
