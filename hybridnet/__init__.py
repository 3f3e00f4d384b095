"""Import shim: the product package lives in the directory the build contract names
(``transformer-cnn-hybrid-network-for-video-processing_b200/``), which is not a valid Python
identifier; ``import hybridnet`` maps onto it."""
import os as _os

_PKG_DIR = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))),
                         "transformer-cnn-hybrid-network-for-video-processing_b200")
__path__.append(_PKG_DIR)

from ._pkg_init import *  # noqa: E402,F401,F403
from ._pkg_init import MODELS  # noqa: E402,F401
