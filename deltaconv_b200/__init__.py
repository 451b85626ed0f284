"""deltaconv_b200 — B200-native (sm_100a) implementation of DeltaConv's data-parallel hot path
behind the reference's Python API (rubenwiersma/deltaconv 1.0.2)."""
__version__ = (0, 1, 0)
