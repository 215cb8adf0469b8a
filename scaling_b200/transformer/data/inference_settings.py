"""Inference-time controls that travel with a batch through the layer IO (reference:
transformer/data/inference_settings.py): KV-cache handling, which layers' hidden states to export as embeddings,
where image tokens sit, and the attention-manipulation ("AtMan") controls per batch item."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import List, NamedTuple, Optional, Tuple, Union

from pydantic import Field

from ...core import BaseConfig


class Control(BaseConfig):
    """Scale the attention every query pays to one token."""

    token_index: int = Field(description="token index to be controlled")
    factor: float = Field(description="control factor")


class InferenceSuppressionParameters(NamedTuple):
    """Controls of ONE batch item; tokens similar to a controlled token above the threshold are controlled too."""

    contextual_control_threshold: Optional[float]
    control_log_additive: bool
    controls: Optional[List[Control]]


def _log_additive_mode(params: Optional[List[InferenceSuppressionParameters]]) -> Union[bool, List[bool]]:
    """One flag when the whole batch agrees (the common case — lets the attention add / multiply the manipulation
    matrix without a per-item branch), the per-item list otherwise."""
    flags = [p.control_log_additive for p in (params or [])]
    if not flags or all(flags):
        return True
    return False if not any(flags) else flags


@dataclass
class InferenceSettings:
    use_cache: bool
    reset_cache: bool
    cache_index: int
    embedding_layers: List[int]
    input_image_locations: Optional[List[Tuple[int, int, int]]] = None
    inference_control_parameters: Optional[List[InferenceSuppressionParameters]] = None
    control_log_additive_batch: Union[bool, List[bool]] = field(init=False)

    def __post_init__(self) -> None:
        self.control_log_additive_batch = _log_additive_mode(self.inference_control_parameters)
