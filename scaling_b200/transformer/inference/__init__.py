from .inference_model import CompletionOutput, TransformerInferenceModule  # noqa: F401
from .sample import (  # noqa: F401
    fast_multinomial,
    sample_argmax,
    sample_temperature,
    sample_top_k,
    sample_top_p,
    top_k_transform,
    top_p_transform,
)
