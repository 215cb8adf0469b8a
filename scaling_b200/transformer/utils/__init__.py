from .get_tflops import (  # noqa: F401
    HardwareType,
    get_model_flop_utilization_palm,
    get_tflops_aleph_alpha,
    get_tflops_bloom,
    get_tflops_electra,
    get_tflops_megatron,
)
