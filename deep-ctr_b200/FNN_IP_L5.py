"""Entry point mirroring the reference's python/FNN_IP_L5.py: `from deepctr_b200.FNN_IP_L5 import FNN_IP_L5`."""
from .models import FNN_IP_L5  # noqa: F401
