"""Lazy exports, as the reference's models/sota/__init__.py:22-44 does (only the RIFE family is built here)."""


def __getattr__(name):
    if name in ("RIFEModel", "RIFELiteModel", "get_rife_model"):
        from . import rife_wrapper
        return getattr(rife_wrapper, name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
