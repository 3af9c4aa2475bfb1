"""Drop-in name of the reference package (raleng/nmf, nmf/__init__.py:1): `from nmf import NMF` and
`from nmf.mur import mur` / `nmf.admm` / `nmf.ao_admm` / `nmf.utils` resolve to the B200 build (nmf_b200)."""
import sys as _sys

from nmf_b200 import NMF  # noqa: F401
from nmf_b200 import admm as _admm, ao_admm as _ao_admm, mur as _mur, nmf as _nmf, utils as _utils

for _name, _mod in (("mur", _mur), ("admm", _admm), ("ao_admm", _ao_admm), ("utils", _utils), ("nmf", _nmf)):
    _sys.modules[__name__ + "." + _name] = _mod
    globals()[_name] = _mod

__all__ = ["NMF"]
