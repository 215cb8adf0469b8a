"""Model families.  The transformer suite lives in :mod:`scaling_b200.transformer` (same layout as the reference's
``scaling.transformer``); this package is the short import path plus ready-made architecture presets."""
from ..transformer.model import (  # noqa: F401
    TransformerLayerIO,
    TransformerParallelModule,
    get_transformer_layer_specs,
    init_model,
    init_optimizer,
    loss_function,
    metrics_aggregation_fn,
)
from .presets import PRESETS, architecture_preset  # noqa: F401
