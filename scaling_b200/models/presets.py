"""Named ``transformer_architecture`` presets: geometry only — every other field keeps the config default and can be
overridden.  ``mistral_7b`` is the headline benchmark model."""
from __future__ import annotations

from typing import Any, Dict

_LLAMA_STYLE = dict(norm_type="rms", relative_position_embedding_type="rotary_complex", mlp_type="swiglu",
                    attention_bias=False, mlp_bias=False, attention_qkv_in_one=False, weight_tying=False,
                    precision="bfloat16", masked_softmax={"kernel": "flash_attention"})
_GPT_STYLE = dict(norm_type="layernorm", relative_position_embedding_type="rotary", mlp_type="default", mlp_factor=4.0,
                  attention_bias=True, mlp_bias=True, attention_qkv_in_one=True, weight_tying=True,
                  precision="bfloat16", masked_softmax={"kernel": "flash_attention"})

PRESETS: Dict[str, Dict[str, Any]] = {
    # name: vocab, hidden, layers, heads, kv heads, mlp factor (ffn / hidden), sequence length
    "mistral_7b": dict(_LLAMA_STYLE, vocab_size=32000, hidden_size=4096, num_layers=32, num_attention_heads=32,
                       attention_num_kv_heads=8, mlp_factor=3.5, sequence_length=4096),
    "llama2_7b": dict(_LLAMA_STYLE, vocab_size=32000, hidden_size=4096, num_layers=32, num_attention_heads=32,
                      attention_num_kv_heads=None, mlp_factor=2.6875, sequence_length=4096),
    "llama2_13b": dict(_LLAMA_STYLE, vocab_size=32000, hidden_size=5120, num_layers=40, num_attention_heads=40,
                       attention_num_kv_heads=None, mlp_factor=2.7, sequence_length=4096),
    "llama2_70b": dict(_LLAMA_STYLE, vocab_size=32000, hidden_size=8192, num_layers=80, num_attention_heads=64,
                       attention_num_kv_heads=8, mlp_factor=3.5, sequence_length=4096),
    "llama3_8b": dict(_LLAMA_STYLE, vocab_size=128256, hidden_size=4096, num_layers=32, num_attention_heads=32,
                      attention_num_kv_heads=8, mlp_factor=3.5, sequence_length=8192, rotary_embedding_base=500000),
    "gpt_1b3": dict(_GPT_STYLE, vocab_size=50304, hidden_size=2048, num_layers=24, num_attention_heads=16,
                    sequence_length=2048),
    "gpt_6b7": dict(_GPT_STYLE, vocab_size=50304, hidden_size=4096, num_layers=32, num_attention_heads=32,
                    sequence_length=2048),
    "tiny": dict(_LLAMA_STYLE, vocab_size=512, hidden_size=256, num_layers=2, num_attention_heads=2,
                 attention_num_kv_heads=1, mlp_factor=2.0, sequence_length=128),
}


def architecture_preset(name: str, **overrides: Any) -> Dict[str, Any]:
    """A ``transformer_architecture`` dict for ``TransformerConfig.from_dict({... "transformer_architecture": ...})``."""
    if name not in PRESETS:
        raise KeyError(f"unknown preset {name!r}; available: {sorted(PRESETS)}")
    arch = {k: (dict(v) if isinstance(v, dict) else v) for k, v in PRESETS[name].items()}
    arch.update(overrides)
    return arch
