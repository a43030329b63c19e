"""Hand-written sm_100a operators and their Python bindings."""
from . import _lib

__all__ = ['_lib']
