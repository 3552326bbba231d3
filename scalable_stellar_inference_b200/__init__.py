"""B200-native NUTS potential for celestify's stellar models (alexlyttle/scalable-stellar-inference).

Python surface mirrors the reference package ``celestify``:

    Emulator(backend="cuda")                      src/celestify/emulator.py:39-51
    BolometricCorrections(bands, backend="cuda")  src/celestify/bolometric_corrections.py:54-60
    Star(bands, backend="cuda")                   src/celestify/star.py:33-39
    numpyro_models.SingleStarModel / MultiStarModel / HierarchicalStarModel
                                                  src/celestify/numpyro_models.py:41,248,262

All arithmetic of the hot path runs in libcsi.so (hand-written sm_100a CUDA, C ABI in include/csi.h).
"""
import os

PACKAGEDIR = os.path.abspath(os.path.dirname(__file__))
__version__ = "0.1.0"

from .emulator import Emulator  # noqa: E402,F401
from .bolometric_corrections import BolometricCorrections  # noqa: E402,F401
from .star import Star  # noqa: E402,F401
from . import numpyro_models  # noqa: E402,F401
from .numpyro_models import SingleStarModel, MultiStarModel, HierarchicalStarModel  # noqa: E402,F401
