"""Transformer suite — same names as ``scaling.transformer`` (reference transformer/__init__.py:3-14)."""
from .config import TransformerConfig  # noqa: F401
from .context import TransformerContext  # noqa: F401
from .data import TextBlendedDataset, TextDataset, TextDatasetBatch, TextDatasetItem  # noqa: F401
from .model import (  # noqa: F401
    TransformerLayerIO,
    TransformerParallelModule,
    get_transformer_layer_specs,
    init_model,
    init_optimizer,
    loss_function,
    metrics_aggregation_fn,
)


def __getattr__(name: str):
    """Lazy imports for the heavier dataset families."""
    if name in ("FinetuningTextDataset", "FinetuningTextBlendedDataset", "FinetuningChatDataset", "FinetuningChatBlendedDataset"):
        from .data import finetuning

        return getattr(finetuning, name)
    if name == "LegacyBlendedDataset":
        from .data.legacy import LegacyBlendedDataset

        return LegacyBlendedDataset
    raise AttributeError(name)
