"""scaling_b200 — a Blackwell-native (sm_100a) 3D-parallel training framework with the capabilities and public API of
Aleph Alpha's *Scaling* (``scaling.core`` -> ``scaling_b200.core``, ``scaling.transformer`` -> ``scaling_b200.transformer``)."""

__version__ = "0.1.0"
