"""B200-native hot path of Atomu2014/deep-ctr (import name: `deepctr_b200`; directory `deep-ctr_b200/`).

Kernels: csrc/*.cu -> libdctr.so (C-ABI in include/dctr.h).  Host side: engine.py (device state + step
orchestration), models.py / theano_fnn.py / snn.py (the reference's model interface), data.py (formats + synthetic data),
ingest.py (device-side parsing of the text formats), dist.py (NCCL data-parallel and row-sharded steps).
"""
from . import _lib
from ._lib import build, launch_count

__all__ = ["_lib", "build", "launch_count"]


def __getattr__(name):
    # torch-dependent modules are imported lazily so `import deepctr_b200` and the ABI tests stay light
    if name in ("FM", "LR", "FNN", "FNN_IP_L3", "FNN_IP_L5", "FNN_IP_L7", "FNN_OP_L3", "Session", "init_var_map"):
        from . import models
        return getattr(models, name)
    if name == "FNN_wnzh":
        from .theano_fnn import FNN_wnzh
        return FNN_wnzh
    if name == "SNN":
        from .snn import SNN
        return SNN
    if name == "DeviceText":
        from .ingest import DeviceText
        return DeviceText
    if name in ("engine", "models", "theano_fnn", "snn", "ingest", "metrics", "data", "dist"):
        import importlib
        return importlib.import_module("." + name, __name__)
    raise AttributeError(name)
