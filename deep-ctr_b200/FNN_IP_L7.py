"""Entry point mirroring the reference's python/FNN_IP_L7.py: `from deepctr_b200.FNN_IP_L7 import FNN_IP_L7`."""
from .models import FNN_IP_L7  # noqa: F401
