"""Import shim: the product package lives in ``viscoin-next_b200/`` (a directory name Python cannot import
directly).  This module makes it importable as ``viscoin_next_b200`` by pointing the package path there."""

import os as _os

_real = _os.path.join(_os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))), "viscoin-next_b200")
__path__ = [_real]
with open(_os.path.join(_real, "__init__.py")) as _fh:
    exec(compile(_fh.read(), _os.path.join(_real, "__init__.py"), "exec"))
del _os, _fh, _real
