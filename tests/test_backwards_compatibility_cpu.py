"""Forward pass of a stored legacy checkpoint against its stored per-sub-module activations.

The fixture (weights of a 1-layer model from the code base that preceded the reference, the activations that code base
produced, and the key translation) belongs to the reference checkout
(tests/transformer/files/backward_compatibility_checkpoint, used by its tests/transformer/test_backwards_compatibility.py:132-237);
it is read in place and the test is skipped where that checkout is not present.  Tolerance as in the reference: 3e-3."""
from pathlib import Path

import pytest
import torch

from tests.utils import dist_launcher

FIXTURE = Path("/root/reference/tests/transformer/files/backward_compatibility_checkpoint")


def _translate(legacy):
    """Key names of the legacy code base -> layer-indexed names (embedding 0, block 1, final norm 2, tied head 3)."""
    out = {}
    for k, v in legacy.items():
        if k.endswith(".inv_freq"):
            continue
        if k == "transformer.embeddings.word_embeddings.weight":
            out["_layers.0.embedding.weight"] = v
            out["_layers.3.embedding.weight"] = v
        elif k.startswith("transformer.layer0."):
            k2 = k.replace("transformer.layer0.", "_layers.1.").replace(".attention.", ".self_attention.")
            out[k2.replace("dense_h_to_4h", "dense_in").replace("dense_4h_to_h", "dense_out")] = v
        elif k.startswith("transformer.norm."):
            out[k.replace("transformer.", "_layers.2.")] = v
        else:
            raise KeyError(k)
    return out


def _worker(rank, world_size, master_port):
    from scaling_b200.core import Topology
    from scaling_b200.transformer import TransformerConfig, TransformerContext, init_model
    from scaling_b200.transformer.data.text_dataset_batch import TextDatasetBatch
    from scaling_b200.transformer.data.utils import get_cumulative_seq_lengths, get_position_ids

    config = TransformerConfig.from_dict({
        "topology": {"world_size": 1, "global_rank": 0, "local_slot": 0, "model_parallel_size": 1, "data_parallel_size": 1,
                     "micro_batch_size": 2, "gradient_accumulation_steps": 1},
        "transformer_architecture": {"vocab_size": 512, "sequence_length": 4, "hidden_size": 16, "num_attention_heads": 2,
                                     "num_layers": 1},
    })
    topology = Topology(config.topology)
    context = TransformerContext(config=config, topology=topology)
    context.initialize(master_addr="127.0.0.1", master_port=str(master_port), seed=config.trainer.seed)
    model = init_model(context=context)
    model.eval()
    missing, unexpected = model.load_state_dict(_translate(torch.load(FIXTURE / "state_dict.pt", weights_only=False)),
                                                strict=False)
    assert not unexpected and not [m for m in missing if "rotary" not in m and "cos" not in m and "sin" not in m], (missing, unexpected)
    truth = torch.load(FIXTURE / "ground_truth.pt", weights_only=False)
    ids = truth["input"].to(topology.device)
    batch = TextDatasetBatch(input_token_ids=ids,
                             cumulative_seq_lengths=get_cumulative_seq_lengths(ids, reset_attention_mask=False),
                             position_ids=get_position_ids(ids, reset_position_ids=False))
    layers = model._layers
    with torch.no_grad():
        emb = layers[0](batch)
        block = layers[1](emb)
        got = {
            "hidden_states_embedding": emb.activations,
            "hidden_states_input_layernorm": layers[1].input_layernorm(emb.activations),
            "hidden_states_layer0": block.activations,
        }
        attn = layers[1].attention_block(emb.activations, cumulative_seq_lengths=batch.cumulative_seq_lengths,
                                         position_ids=batch.position_ids)
        got["hidden_states_attention"] = attn
        got["hidden_states_post_attention_layernorm"] = layers[1].post_attention_layernorm(attn)
        got["hidden_states_mlp"] = layers[1].mlp_block(attn)
        norm = layers[2](block)
        got["hidden_states_norm"] = norm.activations
        got["output_logits"] = layers[3](norm).activations
    return {k: float((v.float().cpu() - truth[k].float()).abs().max()) for k, v in got.items()}


@pytest.mark.skipif(not FIXTURE.is_dir(), reason="reference checkout with the legacy checkpoint fixture is not present")
def test_legacy_checkpoint_reproduces_stored_activations():
    diffs = dist_launcher(_worker, 1)[0]
    assert len(diffs) == 8
    bad = {k: d for k, d in diffs.items() if not d < 3.0e-3}
    assert not bad, f"forward pass differs from the stored activations: {bad} (all: {diffs})"
