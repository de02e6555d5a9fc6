"""Entry point mirroring the reference's python/FM.py: `from deepctr_b200.FM import FM`."""
from .models import FM  # noqa: F401
