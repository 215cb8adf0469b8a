"""NVLink / NVSwitch peer-memory runtime: symmetric workspaces and the fused compute+collective operators built on
them (the torch.distributed/NCCL expressions of the same collectives live in ``core/nn/tp_ops.py``)."""
from .symm import SymmetricArena, TPComm, fused_tp_enabled, fused_zero_enabled, get_tp_comm  # noqa: F401
