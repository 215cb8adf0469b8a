"""Small shared utilities (short import path for the pieces under ``core/utils`` and ``transformer/utils``)."""
from ..core.utils.param_merge import merge_parameter, split_parameter  # noqa: F401
from ..core.utils.port import find_free_port  # noqa: F401
from ..transformer.utils.get_tflops import *  # noqa: F401,F403
