"""Import shim for toolz: the handful of functions gt4py.eve touches."""
from . import functoolz, itertoolz
from .functoolz import complement, curry, identity
