"""Loads the UNMODIFIED reference modules copied to oracle/_ref by oracle/make_ref.py (baseline / test infrastructure).

The files use relative imports (`from .configs import ...`), so they are imported as sub-modules of a synthetic
package whose __path__ is oracle/_ref/models -- the reference's own `models/__init__.py` (which pulls in the whole
model zoo) is never executed, and no file is edited.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = "_deepearth_reference_models"


class _Ref:
    def __init__(self, pkg):
        self.hash_encoding = importlib.import_module(pkg + ".hash_encoding")
        self.configs = importlib.import_module(pkg + ".configs")
        self.encoders = importlib.import_module(pkg + ".encoders")
        self.MultiResolutionHashEncoding = self.hash_encoding.MultiResolutionHashEncoding
        self.Grid4DEncoder = self.encoders.Grid4DEncoder
        self.DeepEarthConfig = self.configs.DeepEarthConfig
        try:
            self.cross_modal_fusion = importlib.import_module(pkg + ".cross_modal_fusion")
            self.SpatialTemporalEmbedding = self.cross_modal_fusion.SpatialTemporalEmbedding
        except Exception:                      # optional: only the f3 comparator needs it
            self.cross_modal_fusion = None
            self.SpatialTemporalEmbedding = None


def load_reference_models():
    """-> object exposing the reference classes, or None when oracle/_ref has not been built."""
    base = os.path.join(_HERE, "_ref", "models")
    if not os.path.exists(os.path.join(base, "hash_encoding.py")):
        return None
    if _PKG not in sys.modules:
        pkg = types.ModuleType(_PKG)
        pkg.__path__ = [base]
        sys.modules[_PKG] = pkg
    return _Ref(_PKG)
