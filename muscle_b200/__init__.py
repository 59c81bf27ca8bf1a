"""muscle_b200 — B200-native (sm_100a) implementation of the multi-strategy contrastive-loss hot
path of SCoulY/MuSCLe, behind the reference's own loss call signatures.

Everything computes in hand-written CUDA (libmuscle_b200.so, C ABI in include/muscle_b200.h);
there is no CPU or eager fallback.
"""
from .ops import dense_infonce, dense_infonce_value_and_grad  # noqa: F401
from .pool import pool_prototypes  # noqa: F401

__all__ = ["dense_infonce", "dense_infonce_value_and_grad", "pool_prototypes"]
