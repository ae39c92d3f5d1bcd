"""k3m_b200 — B200-native K3M pretraining step (sm_100a CUDA kernels behind a C ABI)."""
from .config import BertConfig  # noqa: F401

__all__ = ["BertConfig"]
