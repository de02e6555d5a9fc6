"""Entry point mirroring the reference's python/FNN_OP_L3.py: `from deepctr_b200.FNN_OP_L3 import FNN_OP_L3`."""
from .models import FNN_OP_L3  # noqa: F401
