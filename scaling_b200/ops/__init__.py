"""Native sm_100a operators: C ABI library (``csrc/``), ctypes binding, tensor wrappers and autograd functions."""
from ._lib import available, launch_count, reset_launch_count  # noqa: F401
