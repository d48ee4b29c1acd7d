"""Import alias: the package directory is ``evi-rag_b200/`` (not a Python identifier).

``import evi_rag_b200`` resolves here and forwards to the real package directory, so every
submodule (``evi_rag_b200.retriever`` ...) is loaded from ``evi-rag_b200/``.
"""
import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "evi-rag_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py"), "r", encoding="utf-8") as _f:
    exec(compile(_f.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _os, _f, _real
