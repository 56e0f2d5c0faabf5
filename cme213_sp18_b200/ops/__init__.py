"""Op layer: `host` = native C++/OpenMP (CPU), `device` = hand-written sm_100a CUDA kernels."""
from . import host  # noqa: F401

__all__ = ["host", "device"]


def __getattr__(name):
    if name == "device":
        import importlib

        return importlib.import_module(".device", __name__)
    raise AttributeError(name)
