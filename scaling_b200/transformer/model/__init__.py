from .layers import (  # noqa: F401
    EmbeddingInput,
    LayerNormWrapper,
    TransformerEmbeddingHead,
    TransformerLayer,
    TransformerLayerBaseIO,
    TransformerLayerIO,
    TransformerLMHead,
    TransformerLMHeadTied,
)
from .model import (  # noqa: F401
    TransformerParallelModule,
    get_parameter_groups,
    get_transformer_layer_specs,
    init_model,
    init_optimizer,
    loss_function,
    metrics_aggregation_fn,
)
