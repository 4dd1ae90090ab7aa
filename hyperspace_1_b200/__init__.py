"""Import shim: the package directory the layout asks for is `hyperspace-1_b200/`,
which is not a valid Python identifier. This package forwards to it."""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "hyperspace-1_b200")
__path__.insert(0, _real)

from . import abi  # noqa: E402,F401

from .index_config import IndexConfig, IllegalArgumentException  # noqa: E402,F401
from .dataframe import Session, Col  # noqa: E402,F401
from .hyperspace import Hyperspace, HyperspaceException, NoChangesException  # noqa: E402,F401
