"""mantrap_b200: the B200-native planning hot path of simon-schaefer/mantrap behind the reference's own interface.

    import mantrap_b200 as mantrap
    mantrap.agents / mantrap.environment / mantrap.modules / mantrap.solver / mantrap.constants

Sub-modules are imported lazily: importing the package needs neither a GPU nor the built library (using it does).
"""
import importlib

_SUBMODULES = ("agents", "api", "attention", "autograd", "build", "constants", "engine", "environment", "modules", "mpc", "parallel", "runtime",
               "scenarios", "solver")


def __getattr__(name):
    if name in _SUBMODULES:
        return importlib.import_module(f"{__name__}.{name}")
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")
