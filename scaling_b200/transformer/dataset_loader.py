"""Dataset selection from ``DataConfig`` flags (reference: transformer/dataset_loader.py:1-136)."""
from __future__ import annotations

from pathlib import Path
from typing import Any, List, Optional, Sequence, Tuple

from ..core import BaseDataset
from .config import DataConfig, TransformerArchitectureConfig, TransformerConfig
from .data.text_dataset import TextDataset


def _softprompt_tokens(a: TransformerArchitectureConfig) -> int:
    return 0 if a.softprompt_config is None else a.softprompt_config.n_tokens


def _tokenizers(a: TransformerArchitectureConfig):
    from .tokenizer import load_tokenizers

    assert a.vocab_file is not None, "vocab_file needs to be set to load the vocabulary file."
    return load_tokenizers(a.vocab_file)


def _build(prefixes: Sequence[Path], data: DataConfig, a: TransformerArchitectureConfig, config: TransformerConfig) -> List[BaseDataset]:
    seed = config.trainer.seed
    if data.finetuning_dataset:
        from .data.finetuning import FinetuningTextDataset

        assert data.use_mmap, "Finetuning is currently only supported with use_mmap set to true."
        tok, tok_nps = _tokenizers(a)
        return [FinetuningTextDataset(data_prefix=p, sequence_length=a.sequence_length, seed=seed,
                                      softprompt_n_tokens=_softprompt_tokens(a), tokenizer=tok,
                                      tokenizer_no_prefix_space=tok_nps, memory_map_dataset=data.finetuning_dataset_memory_map)
                for p in prefixes]
    if data.finetuning_chat_dataset:
        from .data.finetuning import FinetuningChatDataset

        tok, tok_nps = _tokenizers(a)
        return [FinetuningChatDataset(data_path=p, sequence_length=a.sequence_length, seed=seed,
                                      softprompt_n_tokens=_softprompt_tokens(a), tokenizer=tok,
                                      tokenizer_no_prefix_space=tok_nps) for p in prefixes]
    return [TextDataset(data_prefix=p, sequence_length=a.sequence_length, seed=seed, legacy_dataset=data.legacy_dataset,
                        load_mmap_index_to_memory=data.load_mmap_index_to_memory,
                        load_data_item_mmap_index_to_memory=data.load_data_item_mmap_index_to_memory,
                        only_full_sequences=data.only_full_sequences,
                        allow_incomplete_sequences_every_n=data.allow_incomplete_sequences_every_n, use_mmap=data.use_mmap)
            for p in prefixes]


def load_datasets(data_config: DataConfig, architecture_config: TransformerArchitectureConfig,
                  config: TransformerConfig) -> Tuple[List[BaseDataset], List[BaseDataset]]:
    assert data_config.data_prefixes is not None, "path to data prefix not defined in Transformer context"
    train = _build(data_config.data_prefixes, data_config, architecture_config, config)
    val = _build(data_config.validation_data_prefixes, data_config, architecture_config, config) \
        if data_config.validation_data_prefixes else []
    return train, val
