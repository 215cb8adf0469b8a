"""Text generation on top of the single-process inference pipeline (reference:
transformer/inference/inference_model.py:1-263).

``from_checkpoint`` reads ``config.yml`` + ``vocab.json`` + the layer-wise ``model_state_layer_*.pt`` files a trainer
wrote (any training layout — the files are TP-merged and keyed by global layer index).  ``generate`` decodes greedily
or with a sampler, either re-running the growing sequence (uncached) or feeding one token at a time against the
attention layers' KV cache (cached, batch size 1).
"""
from __future__ import annotations

from pathlib import Path
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import torch
import yaml

from ...core.nn.inference_module import InferenceModule, RecorderSetting
from ...core.nn.layers import LayerSpec
from ...core.topology import PipePartitionMethod
from ..config import TransformerArchitectureConfig
from ..data.inference_settings import InferenceSettings
from ..data.text_dataset_batch import TextDatasetBatch
from ..model.layers import TransformerLayerIO
from ..model.model import get_transformer_layer_specs
from ..tokenizer import Tokenizer
from .sample import sample_argmax

SampleFn = Callable[[torch.Tensor], torch.Tensor]


class CompletionOutput:
    def __init__(self, completion_text: Optional[str], completion_tokens: List[int], completion_logits: torch.Tensor):
        self.completion_text = completion_text
        self.completion_tokens = completion_tokens
        self.completion_logits = completion_logits


class TransformerInferenceModule(InferenceModule):
    def __init__(self, layer_specs: List[LayerSpec], devices: Optional[Sequence[Any]] = None,
                 pipe_partition_method: PipePartitionMethod = PipePartitionMethod.UNIFORM,
                 pipe_partition_overwrite: Optional[List[int]] = None, tokenizer: Optional[Tokenizer] = None):
        super().__init__(layer_specs=layer_specs, devices=devices, pipe_partition_method=pipe_partition_method,
                         pipe_partition_overwrite=pipe_partition_overwrite)
        self.tokenizer = tokenizer

    @staticmethod
    def _parse_config_file(config_file: Path) -> TransformerArchitectureConfig:
        with open(config_file, encoding="utf-8") as f:
            config_dict = yaml.safe_load(f)
        return TransformerArchitectureConfig.from_dict(config_dict["transformer_architecture"])

    @classmethod
    def from_checkpoint(cls, checkpoint_dir: Path, devices: Optional[Sequence[Any]] = None,
                        pipe_partition_method: PipePartitionMethod = PipePartitionMethod.UNIFORM,
                        pipe_partition_overwrite: Optional[List[int]] = None, config_file: Optional[Path] = None,
                        vocab_file: Optional[Path] = None) -> "TransformerInferenceModule":
        checkpoint_dir = Path(checkpoint_dir)
        config_file = Path(config_file) if config_file is not None else checkpoint_dir / "config.yml"
        vocab_file = Path(vocab_file) if vocab_file is not None else checkpoint_dir / "vocab.json"
        assert config_file.is_file(), "Config file not found"
        assert vocab_file.is_file(), "Vocab file not found"
        architecture = cls._parse_config_file(config_file)
        model = cls(layer_specs=get_transformer_layer_specs(architecture_config=architecture), devices=devices,
                    pipe_partition_method=pipe_partition_method, pipe_partition_overwrite=pipe_partition_overwrite,
                    tokenizer=Tokenizer.from_file(str(vocab_file)))
        model.load_checkpoint(checkpoint_dir)
        return model

    # ---- one forward ------------------------------------------------------------------------------------------
    def _pre_process_input(self, input_text: Optional[str] = None, input_tokens: Optional[List[int]] = None,
                           process_for_cached_inference: bool = True) -> TextDatasetBatch:
        assert (input_text is None) ^ (input_tokens is None), "Either input_text or input_tokens needs to be provided"
        if input_text is not None:
            assert self.tokenizer is not None
            input_tokens = self.tokenizer.encode(input_text)
        settings = InferenceSettings(use_cache=process_for_cached_inference, reset_cache=True, cache_index=0,
                                     embedding_layers=[-1])
        return TextDatasetBatch(input_token_ids=torch.tensor(input_tokens).unsqueeze(0), inference_settings=settings)

    @staticmethod
    def _post_process_output(output: TransformerLayerIO) -> torch.Tensor:
        return output.activations.squeeze()

    def logits(self, input_text: Optional[str] = None, input_tokens: Optional[List[int]] = None) -> torch.Tensor:
        """Final activations (logits) [seq, vocab] for one prompt."""
        out = self.forward(self._pre_process_input(input_text=input_text, input_tokens=input_tokens))
        return self._post_process_output(out).squeeze()

    def logits_with_hidden_state_recorder(self, input_text: Optional[str] = None, input_tokens: Optional[List[int]] = None,
                                          recorder_settings_per_layer: Optional[Dict[int, RecorderSetting]] = None
                                          ) -> Tuple[torch.Tensor, Dict[int, Dict[str, Any]]]:
        batch = self._pre_process_input(input_text=input_text, input_tokens=input_tokens)
        out, record = self.forward_with_hidden_state_recorder(batch, recorder_settings_per_layer=recorder_settings_per_layer)
        return self._post_process_output(out), record

    # ---- decoding ---------------------------------------------------------------------------------------------
    def generate(self, max_tokens: int, input_text: Optional[str] = None, input_tokens: Optional[List[int]] = None,
                 sample_fn: SampleFn = sample_argmax, stop_tokens: Optional[Sequence[int]] = None,
                 use_cache: bool = True) -> CompletionOutput:
        if stop_tokens is None:
            assert self.tokenizer is not None, "If no tokenizer is provided, a stop token needs to be set manually"
            stop_tokens = [self.tokenizer.eos_token_id]
        batch = self._pre_process_input(input_text=input_text, input_tokens=input_tokens,
                                        process_for_cached_inference=use_cache)
        prompt_len = batch.input_token_ids.shape[-1]
        step_settings = InferenceSettings(use_cache=use_cache, reset_cache=not use_cache, cache_index=0, embedding_layers=[-1])
        tokens: List[int] = []
        step_logits: List[torch.Tensor] = []
        out: Optional[TransformerLayerIO] = None
        for k in range(max_tokens):
            out = self.forward(batch)
            nxt = sample_fn(out.activations).reshape(1)
            tok = int(nxt.item())
            tokens.append(tok)
            if use_cache:  # feed just the new token; the KV cache holds the past
                step_logits.append(out.activations[:, -1, :])
                batch = TextDatasetBatch(input_token_ids=nxt.reshape(1, 1).cpu(), inference_settings=step_settings,
                                         position_ids=torch.tensor([[prompt_len + k]]))
            else:  # re-run the whole sequence
                ids = torch.cat([batch.input_token_ids.cpu(), nxt.reshape(1, 1).cpu()], dim=-1)
                batch = TextDatasetBatch(input_token_ids=ids, inference_settings=step_settings)
            if tok in stop_tokens:
                break
        if use_cache:
            logits = torch.cat(step_logits) if step_logits else torch.empty(0)
        else:
            logits = self._post_process_output(out)[prompt_len - 1:] if out is not None else torch.empty(0)
        text = self.tokenizer.decode(tokens) if self.tokenizer is not None else None
        return CompletionOutput(completion_text=text, completion_tokens=tokens, completion_logits=logits)
