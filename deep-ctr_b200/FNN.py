"""Entry point mirroring the reference's python/FNN.py: `from deepctr_b200.FNN import FNN`."""
from .models import FNN  # noqa: F401
