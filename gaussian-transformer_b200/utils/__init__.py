"""`utils` of the drop-in: only `loss_utils` lives here (fused L1 + SSIM, `utils/loss_utils.py:17-63` of the reference).

The reference's own `utils/` directory has no `__init__.py` (a namespace package), so this regular package wins the
import wherever it sits on `sys.path` — and would hide `utils.system_utils`, `utils.general_utils`, `utils.sh_utils`,
`utils.graphics_utils`, ... that `scene/__init__.py:15`, `scene/gaussian_model.py:14-21` and `train.py:14-22` import.
The package path is therefore extended with every other `utils` directory found along `sys.path` (and the working
directory, which is where `python train.py` runs): `utils.loss_utils` resolves here first, every other submodule falls
through to the reference's file, unmodified.
"""
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
try:
    _cwd = os.getcwd()
except OSError:  # the working directory was removed under the process
    _cwd = None
_cand = None
for _root in [""] + list(sys.path):
    if not (_root or _cwd):
        continue
    _cand = os.path.abspath(os.path.join(_root or _cwd, "utils"))
    if _cand != _here and os.path.isdir(_cand) and _cand not in __path__:
        __path__.append(_cand)
del _root, _cand, _cwd
