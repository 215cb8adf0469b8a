"""Typed run context of the transformer suite (reference: transformer/context/context.py)."""
from ..core import DeterminedBaseContext
from .config import TransformerConfig


class TransformerContext(DeterminedBaseContext[TransformerConfig]):
    config: TransformerConfig
