"""Import shim: the package directory is `deep-ctr_b200/` (not a legal Python identifier), so this stub
points `deepctr_b200`'s search path there and executes its __init__."""
import os

_real = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "deep-ctr_b200")
__path__ = [_real]
with open(os.path.join(_real, "__init__.py")) as _f:
    exec(compile(_f.read(), os.path.join(_real, "__init__.py"), "exec"))
