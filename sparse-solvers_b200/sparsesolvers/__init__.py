"""sparsesolvers -- B200-native stand-in for the reference's Python package.

The reference wheel exposes `sparsesolvers.Homotopy`, `sparsesolvers.Irls` and `version()` from its
pybind11 module (bindings/python/sparsesolvers/binding.cpp:114-149, README.md:15-28).  The same names,
argument meaning and error behaviour are provided here on top of the C ABI in include/ssb200.h.
"""
from .binding import *  # noqa: F401,F403
from .binding import __all__  # noqa: F401
