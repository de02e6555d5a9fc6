"""Entry point mirroring the reference's python/LR.py: `from deepctr_b200.LR import LR`."""
from .models import LR  # noqa: F401
