"""Importable alias of the product package.

The product lives in ``ubisoft-laforge-daft-exprt_b200/`` (the directory name the
build contract asks for).  That name has hyphens and cannot be imported, so this
package points its ``__path__`` at it: ``import daft_exprt_b200.model`` loads
``ubisoft-laforge-daft-exprt_b200/model.py``.
"""
import os as _os

_HERE = _os.path.dirname(_os.path.abspath(__file__))
PACKAGE_DIR = _os.path.join(_os.path.dirname(_HERE), "ubisoft-laforge-daft-exprt_b200")
if not _os.path.isdir(PACKAGE_DIR):  # pragma: no cover
    raise ImportError(f"product directory missing: {PACKAGE_DIR}")
__path__.append(PACKAGE_DIR)

__version__ = "0.1.0"
