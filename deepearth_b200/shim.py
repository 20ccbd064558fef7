"""Import shim: lets reference scripts keep their own import lines.

The reference's callers do ``from models.hash_encoding import MultiResolutionHashEncoding``,
``from models.encoders import Grid4DEncoder`` and ``from models.configs import DeepEarthConfig``
(/root/reference/examples/quick_test.py:15-19, models/__init__.py:13-18).  ``install()`` registers
module objects under those names that resolve to the B200 classes, without touching the reference tree.
"""
from __future__ import annotations

import sys
import types

_NAMES = {
    "models.hash_encoding": ("MultiResolutionHashEncoding",),
    "models.encoders": ("Grid4DEncoder",),
    "models.configs": ("DeepEarthConfig", "ModalityConfig", "TransformerConfig"),
}


def install(package: str = "models", force: bool = False) -> None:
    """Register ``<package>.hash_encoding / .encoders / .configs`` in ``sys.modules``.

    With ``force=False`` an already-imported reference module of the same name is left alone (so the
    reference and this build can be compared in one process); ``force=True`` overrides it.
    """
    import deepearth_b200 as de

    root = sys.modules.get(package)
    if root is None:
        root = types.ModuleType(package)
        root.__path__ = []  # mark as a package
        sys.modules[package] = root
    for full, attrs in _NAMES.items():
        name = full.replace("models", package, 1)
        if name in sys.modules and not force:
            continue
        mod = types.ModuleType(name)
        mod.__doc__ = f"deepearth_b200 shim for the reference module {full}"
        for a in attrs:
            setattr(mod, a, getattr(de, a))
        sys.modules[name] = mod
        setattr(root, name.rsplit(".", 1)[1], mod)
    for attrs in _NAMES.values():
        for a in attrs:
            if force or not hasattr(root, a):
                setattr(root, a, getattr(de, a))
