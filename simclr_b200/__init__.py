"""B200-native NT-Xent (SimCLR contrastive loss): drop-in for nzw0301/SimCLR's loss path."""
from .loss import NT_Xent, NTXentFunction  # noqa: F401
from . import distributed_utils  # noqa: F401

__all__ = ["NT_Xent", "NTXentFunction", "distributed_utils"]
