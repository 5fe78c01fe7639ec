"""explainn_b200 -- B200-native (sm_100a) implementation of the ExplaiNN hot path.

Drop-in module names of the reference package (`explainn/__init__.py`):
    from explainn_b200 import networks, train, test, interpretation
"""
from . import _cabi  # noqa: F401  (fails loudly if the CUDA library is not built)
from . import networks, train, test, interpretation, predict, tools  # noqa: F401

__version__ = "0.1.0"
