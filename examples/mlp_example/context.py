from scaling_b200.core import BaseContext, Topology

from .config import MLPConfig


class MLPContext(BaseContext[MLPConfig]):
    config: MLPConfig

    def __init__(self, config: MLPConfig, topology: Topology):
        super().__init__(config=config, topology=topology)
